/*
 * prt_oracle.cpp — CPU ORACLE for the path-tracing hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is a plain, scalar CPU restatement of the algorithm the reference
 * (Rickyeeeeee/ParallelRayTracing, /root/reference) runs on its CPU backend, plus the LBVH
 * pieces the reference does not have.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it; the product (libprt_b200.so) never does.
 *
 * Parity status: PINNED — tests/test_oracle_pinned.py checks every function below
 * against the reference's own compiled code (oracle/_ref/libprt_ref.so, built from the
 * sources under /root/reference by oracle/Makefile) and against golden vectors generated from
 * it (tests/golden/, generator tests/golden/make_golden.py).  The reference ships no tests or
 * golden vectors of its own.
 *
 * What each function restates (paths relative to /root/reference):
 *   po_camera_dir      Camera::GetCameraRay              src/core/camera.h:104-132
 *   isect_sphere       Circle::Intersect                 src/core/shape.h:157-203
 *   isect_quad         Quad::Intersect                   src/core/shape.h:213-239
 *   isect_tri          Triangle::Intersect               src/core/shape.h:262-303 (watertight variant)
 *   closest_brute      PrimitiveList::Intersect          src/core/primitive.cpp:21-59
 *   scatter            Lambertian/Metal/Dielectric/Emissive::Scatter, MaterialHandle dispatch
 *                                                        src/core/material.h:11-161, src/core/math.h:10-50
 *   trace_path         CPURenderer::TraceRay             src/backend/cpu/renderer.cpp:59-103 (iterative form,
 *                                                        as cuda_megakernel/renderer.cu:81-119)
 *   po_render          CPURenderer::ProgressiveRender    src/backend/cpu/renderer.cpp:18-57 + Film::AddSample film.cu:37-55
 *   po_resolve_display Film::UpdateDisplay               src/core/film.cu:134-194
 *   po_build_lbvh / closest_bvh : no reference counterpart (the reference brute-forces);
 *                      Morton (30 bit), stable sort, Karras 2012 hierarchy, refit, stack traversal.
 *
 * Build: g++ -O2 -ffp-contract=off (no FMA contraction: every fused op is an explicit fmaf).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <atomic>
#include <thread>
#include <vector>
#include "../include/prt.h"

namespace {

struct V3 { float x, y, z; };
inline V3 mk(float x, float y, float z) { return V3{x, y, z}; }
inline V3 add(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 sub(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 mul(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V3 mulv(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 neg(V3 a) { return mk(-a.x, -a.y, -a.z); }
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V3 normalize(V3 v) { float s = 1.0f / sqrtf(dot(v, v)); return mul(v, s); }
inline float comp(V3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }

/* ---------------------------------------------------------------- RNG (math.h:10-36) */
inline uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x21f0aaadu; x ^= x >> 15; x *= 0x735a2d97u; x ^= x >> 15; return x;
}
inline uint32_t path_key(uint64_t seed, uint32_t pixel, uint32_t sample) {
    uint32_t k = hash32((uint32_t)(seed >> 32) ^ 0x9e3779b9u);
    k = hash32((uint32_t)seed ^ k);
    k = hash32(sample ^ k);
    k = hash32(pixel ^ k);
    return k;
}
struct Rng { uint32_t s; };
inline Rng rng_for_bounce(uint32_t key, uint32_t bounce) { Rng r; r.s = hash32(key + 0x9e3779b9u * (bounce + 1u)); return r; }
inline uint32_t rng_next31(Rng& r) {
    r.s = r.s * 747796405u + 2891336453u;
    uint32_t w = ((r.s >> ((r.s >> 28) + 4u)) ^ r.s) * 277803737u;
    w = (w >> 22) ^ w;
    return w >> 1;
}
/* std::rand() / (RAND_MAX + 1.0f) with rand() := rng_next31 */
inline float rng_uniform(Rng& r) { return (float)(int)rng_next31(r) / 2147483648.0f; }
inline float rng_pm1(Rng& r) { return -1.0f + 2.0f * rng_uniform(r); }
inline V3 rng_unit_vector(Rng& r) {
    for (;;) {
        float x = rng_pm1(r), y = rng_pm1(r), z = rng_pm1(r);
        V3 p = mk(x, y, z);
        float l2 = dot(p, p);
        if (1e-8f < l2 && l2 <= 1.0f) { float l = sqrtf(l2); return mk(p.x / l, p.y / l, p.z / l); }
    }
}

/* ---------------------------------------------------------------- scene */
struct Tri { V3 v[3]; V3 n[3]; int mat; };
struct Scene {
    std::vector<prt_sphere> spheres;
    std::vector<prt_quad> quads;
    std::vector<Tri> tris;
    std::vector<prt_material> mats;
    int tri_prim_base = 0;
    float sky[3] = {0.4f, 0.3f, 0.6f};
    /* LBVH (input order q: spheres, quads, triangles) */
    std::vector<V3> amin, amax;
    float bounds[6];
    std::vector<uint32_t> morton, sorted_prim, sorted_key;
    std::vector<int> children, parent, leaf_parent;
    std::vector<V3> node_min, node_max;
    struct WNode { float org[3], scale[3]; int code[4]; uint32_t qlo[3], qhi[3]; };   /* = device WideNode, 64 bytes of content */
    std::vector<WNode> wide;            /* indexed like the internal nodes; only even-depth entries are meaningful */
    std::vector<uint8_t> wide_valid;
    bool built = false;
    int ploc_rounds = 0, height = 0, builder_used = 0;
    int builder = 1, ploc_radius = 4;   /* 0: Karras hierarchy over the sorted codes (LBVH); 1: PLOC over the same sorted order */
    int rr_start = -1; float clamp = INFINITY;      /* estimator options of the device path (default off), po_set_estimator */
    int n_prims() const { return (int)(spheres.size() + quads.size() + tris.size()); }
};

struct Hit { float t; int q; int prim_id; float u, v; };   /* q = input-order index, -1 = miss */
inline void hit_init(Hit& h) { h.t = INFINITY; h.q = -1; h.prim_id = 0x7fffffff; h.u = h.v = 0.f; }
/* strict < on distance, first list entry wins ties (primitive.cpp:44) */
inline void hit_update(Hit& h, float t, int q, int prim_id, float u, float v) {
    if (t < h.t || (t == h.t && prim_id < h.prim_id)) { h.t = t; h.q = q; h.prim_id = prim_id; h.u = u; h.v = v; }
}

/* Circle::Intersect, shape.h:157-203; l = InvMat*o for a translation, d = normalized direction */
float isect_sphere(const prt_sphere& s, V3 o, V3 d, bool& front) {
    V3 l = sub(o, mk(s.cx, s.cy, s.cz));
    float a = dot(d, d);
    float b = 2.0f * dot(l, d);
    float c = dot(l, l) - s.radius * s.radius;
    float disc = b * b - 4.0f * a * c;
    if (!(disc >= 0.0f)) return -1.0f;
    float sq = sqrtf(disc);
    float t1 = (-b + sq) / (2.0f * a);
    float t2 = (-b - sq) / (2.0f * a);
    if (t1 >= s.tmin && t2 >= s.tmin) { front = true; return t1 < t2 ? t1 : t2; }
    front = false;
    if (t1 >= s.tmin) return t1;
    if (t2 >= s.tmin) return t2;
    return -1.0f;
}

/* Quad::Intersect, shape.h:213-239, in world space */
float isect_quad(const prt_quad& q, V3 o, V3 d, bool& front) {
    V3 n = mk(q.nx, q.ny, q.nz);
    V3 oc = sub(o, mk(q.cx, q.cy, q.cz));
    float dn = dot(d, n);
    if (fabsf(dn) < 1e-8f) return -1.0f;
    float on = dot(oc, n);
    float t = -on / dn;
    V3 p = add(oc, mul(d, t));
    float x = dot(p, mk(q.ux, q.uy, q.uz));
    float z = dot(p, mk(q.vx, q.vy, q.vz));
    if (t > q.tmin && (x * x < q.half_w * q.half_w) && (z * z < q.half_h * q.half_h)) { front = on > 0.0f; return t; }
    return -1.0f;
}

/* watertight ray/triangle (Woop et al. 2013); reference is Moller-Trumbore shape.h:262-303 */
/* kx, ky = the two axes after kz in cyclic order; where the paper swaps them for d[kz] < 0 (winding), the sheared x row is
 * negated instead - the same edge functions q - p, see prt_device.cuh:make_shear - with the flip folded into sx */
struct Shear { int kx, ky, kz; float sx, sy, sz; uint32_t flip; };
inline float flip_sign(float v, uint32_t m) { uint32_t u; memcpy(&u, &v, 4); u ^= m; memcpy(&v, &u, 4); return v; }
Shear make_shear(V3 d) {
    Shear r;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    r.kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    r.kx = r.kz + 1; if (r.kx == 3) r.kx = 0;
    r.ky = r.kx + 1; if (r.ky == 3) r.ky = 0;
    float dz = comp(d, r.kz);
    float dx = comp(d, r.kx), dy = comp(d, r.ky);
    r.sz = 1.0f / dz; r.sx = dx * r.sz; r.sy = dy * r.sz;
    uint32_t sb; memcpy(&sb, &r.sz, 4);
    r.flip = sb & 0x80000000u;
    r.sx = flip_sign(r.sx, r.flip);
    return r;
}
float isect_tri(const Tri& tr, V3 o, const Shear& rs, float tmin, float tmax, float& b1, float& b2) {
    V3 A = sub(tr.v[0], o), B = sub(tr.v[1], o), C = sub(tr.v[2], o);
    float Akz = comp(A, rs.kz), Bkz = comp(B, rs.kz), Ckz = comp(C, rs.kz);
    float Ax = fmaf(-rs.sx, Akz, flip_sign(comp(A, rs.kx), rs.flip)), Ay = fmaf(-rs.sy, Akz, comp(A, rs.ky));
    float Bx = fmaf(-rs.sx, Bkz, flip_sign(comp(B, rs.kx), rs.flip)), By = fmaf(-rs.sy, Bkz, comp(B, rs.ky));
    float Cx = fmaf(-rs.sx, Ckz, flip_sign(comp(C, rs.kx), rs.flip)), Cy = fmaf(-rs.sy, Ckz, comp(C, rs.ky));
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        double CxBy = (double)Cx * (double)By, CyBx = (double)Cy * (double)Bx;
        U = (float)(CxBy - CyBx);
        double AxCy = (double)Ax * (double)Cy, AyCx = (double)Ay * (double)Cx;
        V = (float)(AxCy - AyCx);
        double BxAy = (double)Bx * (double)Ay, ByAx = (double)By * (double)Ax;
        W = (float)(BxAy - ByAx);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return -1.0f;
    float det = (U + V) + W;
    if (det == 0.0f) return -1.0f;
    float Az = rs.sz * Akz, Bz = rs.sz * Bkz, Cz = rs.sz * Ckz;
    float T = fmaf(W, Cz, fmaf(V, Bz, U * Az));
    float rdet = 1.0f / det;
    float t = T * rdet;
    if (!(t >= tmin) || t > tmax) return -1.0f;
    b1 = V * rdet; b2 = W * rdet;
    return t;
}

const float TRI_TMIN = 0.001f;   /* kShapeRayTMin, shape.h:128 */

inline void test_prim(const Scene& sc, int q, V3 o, V3 d, V3 dn, const Shear& rs, Hit& h) {
    int ns = (int)sc.spheres.size(), nq = (int)sc.quads.size();
    if (q < ns) {
        bool front; float t = isect_sphere(sc.spheres[q], o, dn, front);
        if (t >= 0.0f) hit_update(h, t, q, sc.spheres[q].prim_id, front ? 1.0f : 0.0f, 0.0f);
    } else if (q < ns + nq) {
        bool front; float t = isect_quad(sc.quads[q - ns], o, d, front);
        if (t >= 0.0f) hit_update(h, t, q, sc.quads[q - ns].prim_id, front ? 1.0f : 0.0f, 0.0f);
    } else {
        float b1, b2; int ti = q - ns - nq;
        float t = isect_tri(sc.tris[ti], o, rs, TRI_TMIN, h.t, b1, b2);
        if (t >= 0.0f) hit_update(h, t, q, sc.tri_prim_base + ti, b1, b2);
    }
}

/* PrimitiveList::Intersect (primitive.cpp:21-59): loop over every primitive */
Hit closest_brute(const Scene& sc, V3 o, V3 d, uint64_t* n_tests) {
    Hit h; hit_init(h);
    Shear rs = make_shear(d);
    V3 dn = normalize(d);
    int n = sc.n_prims();
    for (int q = 0; q < n; ++q) test_prim(sc, q, o, d, dn, rs, h);
    if (n_tests) *n_tests += (uint64_t)n;
    return h;
}

/* ---------------------------------------------------------------- LBVH */
inline uint32_t expand10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
inline uint32_t quant10(float c, float lo, float hi) {
    float e = hi - lo;
    float x = e > 0.0f ? (c - lo) / e : 0.0f;
    int xi = (int)(x * 1024.0f);
    xi = xi < 0 ? 0 : (xi > 1023 ? 1023 : xi);
    return (uint32_t)xi;
}
inline int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
inline int delta(const std::vector<uint32_t>& k, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint32_t a = k[i], b = k[j];
    if (a == b) return 32 + clz32((uint32_t)i ^ (uint32_t)j);
    return clz32(a ^ b);
}

void prim_bounds(const Scene& sc, int q, V3& lo, V3& hi) {
    int ns = (int)sc.spheres.size(), nq = (int)sc.quads.size();
    if (q < ns) {
        const prt_sphere& s = sc.spheres[q];
        lo = mk(s.cx - s.radius, s.cy - s.radius, s.cz - s.radius);
        hi = mk(s.cx + s.radius, s.cy + s.radius, s.cz + s.radius);
    } else if (q < ns + nq) {
        const prt_quad& r = sc.quads[q - ns];
        float c[3] = {r.cx, r.cy, r.cz};
        float a[3] = {r.ux * r.half_w, r.uy * r.half_w, r.uz * r.half_w};
        float b[3] = {r.vx * r.half_h, r.vy * r.half_h, r.vz * r.half_h};
        float l[3], h[3];
        for (int k = 0; k < 3; ++k) {
            float p0 = (c[k] + a[k]) + b[k], p1 = (c[k] + a[k]) - b[k];
            float p2 = (c[k] - a[k]) + b[k], p3 = (c[k] - a[k]) - b[k];
            l[k] = fminf(fminf(p0, p1), fminf(p2, p3));
            h[k] = fmaxf(fmaxf(p0, p1), fmaxf(p2, p3));
        }
        lo = mk(l[0], l[1], l[2]); hi = mk(h[0], h[1], h[2]);
    } else {
        const Tri& t = sc.tris[q - ns - nq];
        lo = mk(fminf(fminf(t.v[0].x, t.v[1].x), t.v[2].x), fminf(fminf(t.v[0].y, t.v[1].y), t.v[2].y), fminf(fminf(t.v[0].z, t.v[1].z), t.v[2].z));
        hi = mk(fmaxf(fmaxf(t.v[0].x, t.v[1].x), t.v[2].x), fmaxf(fmaxf(t.v[0].y, t.v[1].y), t.v[2].y), fmaxf(fmaxf(t.v[0].z, t.v[1].z), t.v[2].z));
    }
}

/* PLOC - parallel locally-ordered clustering (Meister & Bittner 2018) over the Morton-sorted leaves; restatement of
 * lbvh.cu:k_ploc. Every round, each cluster i looks at the clusters within `radius` positions of it in the current
 * (Morton-ordered) array and picks the one whose union box has the smallest half-area (ties: position i^1 if it is among them - so that runs of identical boxes pair up instead of
 * forming a chain -, else the lowest position; either way the globally closest pair is mutual, so every round merges); mutual
 * nearest neighbours merge into a new internal node that takes the place of the lower position; the array is compacted,
 * order kept. Node indices are handed out from the top (n-2) downwards in merge order (by round, then by position), so
 * the last merge - the root - is node 0, like the Karras numbering. Returns false when the clustering does not converge
 * within PLOC_MAX_ROUNDS rounds or the tree is deeper than PLOC_MAX_HEIGHT (the device stack bound): the caller then
 * builds the Karras hierarchy instead, exactly like lbvh.cu. The reference has no counterpart (it has no BVH). */
const int PLOC_MAX_ROUNDS = 192, PLOC_MAX_HEIGHT = 84;
bool build_ploc(Scene& sc) {
    const int n = sc.n_prims(), R = sc.ploc_radius;
    sc.ploc_rounds = 0; sc.height = 0;
    if (n < 2) return true;
    std::vector<int> code(n), code2(n), hgt(n, 0), hgt2(n);  /* code >= 0 internal node, < 0 = ~leaf position */
    std::vector<V3> lo(n), hi(n), lo2(n), hi2(n);
    for (int p = 0; p < n; ++p) { code[p] = ~p; uint32_t q = sc.sorted_prim[p]; lo[p] = sc.amin[q]; hi[p] = sc.amax[q]; }
    std::vector<int> nn(n);
    int c = n, next = n - 1;                               /* the k-th merge from now on gets id next - 1 - k */
    while (c > 1) {
        if (sc.ploc_rounds >= PLOC_MAX_ROUNDS) return false;
        for (int i = 0; i < c; ++i) {
            float best = INFINITY; int bj = -1;
            const int j0 = std::max(0, i - R), j1 = std::min(c - 1, i + R);
            for (int j = j0; j <= j1; ++j) {
                if (j == i) continue;
                const float dx = fmaxf(hi[i].x, hi[j].x) - fminf(lo[i].x, lo[j].x);
                const float dy = fmaxf(hi[i].y, hi[j].y) - fminf(lo[i].y, lo[j].y);
                const float dz = fmaxf(hi[i].z, hi[j].z) - fminf(lo[i].z, lo[j].z);
                const float a = (dx * dy + dy * dz) + dz * dx;
                if (a < best || (a == best && j == (i ^ 1))) { best = a; bj = j; }   /* ties: the partner i^1, else the lowest position */
            }
            nn[i] = bj;
        }
        int out = 0;
        for (int i = 0; i < c; ++i) {
            const int j = nn[i];
            if (nn[j] == i) {
                if (i > j) continue;                       /* absorbed by position j < i */
                const int id = --next;
                sc.children[2 * id] = code[i]; sc.children[2 * id + 1] = code[j];
                if (code[i] >= 0) sc.parent[code[i]] = id; else sc.leaf_parent[~code[i]] = id;
                if (code[j] >= 0) sc.parent[code[j]] = id; else sc.leaf_parent[~code[j]] = id;
                V3 mn = mk(fminf(lo[i].x, lo[j].x), fminf(lo[i].y, lo[j].y), fminf(lo[i].z, lo[j].z));
                V3 mx = mk(fmaxf(hi[i].x, hi[j].x), fmaxf(hi[i].y, hi[j].y), fmaxf(hi[i].z, hi[j].z));
                sc.node_min[id] = mn; sc.node_max[id] = mx;
                code2[out] = id; lo2[out] = mn; hi2[out] = mx; hgt2[out] = std::max(hgt[i], hgt[j]) + 1; ++out;
            } else { code2[out] = code[i]; lo2[out] = lo[i]; hi2[out] = hi[i]; hgt2[out] = hgt[i]; ++out; }
        }
        code.swap(code2); lo.swap(lo2); hi.swap(hi2); hgt.swap(hgt2);
        c = out;
        ++sc.ploc_rounds;
    }
    sc.parent[0] = -1;
    sc.height = hgt[0];
    return sc.height <= PLOC_MAX_HEIGHT;
}

void build_lbvh(Scene& sc) {
    int n = sc.n_prims();
    sc.amin.resize(n); sc.amax.resize(n); sc.morton.resize(n); sc.sorted_prim.resize(n); sc.sorted_key.resize(n);
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int q = 0; q < n; ++q) {
        prim_bounds(sc, q, sc.amin[q], sc.amax[q]);
        lo[0] = fminf(lo[0], sc.amin[q].x); lo[1] = fminf(lo[1], sc.amin[q].y); lo[2] = fminf(lo[2], sc.amin[q].z);
        hi[0] = fmaxf(hi[0], sc.amax[q].x); hi[1] = fmaxf(hi[1], sc.amax[q].y); hi[2] = fmaxf(hi[2], sc.amax[q].z);
    }
    for (int k = 0; k < 3; ++k) { sc.bounds[k] = lo[k]; sc.bounds[3 + k] = hi[k]; }
    /* The Morton grid spans the bounds of the box CENTRES, not the scene bounds: one large primitive (a ground quad) no
     * longer spends the 10 bits per axis on empty space (lbvh.cu:k_prim_bounds / k_morton, option morton_centroid = 1). */
    for (int k = 0; k < 3; ++k) { lo[k] = INFINITY; hi[k] = -INFINITY; }
    for (int q = 0; q < n; ++q) {
        float cx = (sc.amin[q].x + sc.amax[q].x) * 0.5f, cy = (sc.amin[q].y + sc.amax[q].y) * 0.5f, cz = (sc.amin[q].z + sc.amax[q].z) * 0.5f;
        lo[0] = fminf(lo[0], cx); lo[1] = fminf(lo[1], cy); lo[2] = fminf(lo[2], cz);
        hi[0] = fmaxf(hi[0], cx); hi[1] = fmaxf(hi[1], cy); hi[2] = fmaxf(hi[2], cz);
    }
    std::vector<std::pair<uint32_t, uint32_t>> kv(n);
    for (int q = 0; q < n; ++q) {
        float cx = (sc.amin[q].x + sc.amax[q].x) * 0.5f, cy = (sc.amin[q].y + sc.amax[q].y) * 0.5f, cz = (sc.amin[q].z + sc.amax[q].z) * 0.5f;
        uint32_t code = (expand10(quant10(cx, lo[0], hi[0])) << 2) | (expand10(quant10(cy, lo[1], hi[1])) << 1) | expand10(quant10(cz, lo[2], hi[2]));
        sc.morton[q] = code; kv[q] = std::make_pair(code, (uint32_t)q);
    }
    std::stable_sort(kv.begin(), kv.end(), [](const std::pair<uint32_t, uint32_t>& a, const std::pair<uint32_t, uint32_t>& b) { return a.first < b.first; });
    for (int i = 0; i < n; ++i) { sc.sorted_key[i] = kv[i].first; sc.sorted_prim[i] = kv[i].second; }
    int ni = n > 1 ? n - 1 : 0;
    sc.children.assign(2 * (size_t)ni, 0); sc.parent.assign(ni, -1); sc.leaf_parent.assign(n, -1);
    sc.node_min.resize(ni); sc.node_max.resize(ni);
    sc.builder_used = 0;
    if (sc.builder == 1 && build_ploc(sc)) { sc.builder_used = 1; sc.built = true; return; }
    sc.children.assign(2 * (size_t)ni, 0); sc.parent.assign(ni, -1); sc.leaf_parent.assign(n, -1);
    const std::vector<uint32_t>& k = sc.sorted_key;
    for (int i = 0; i < ni; ++i) {                       /* Karras 2012, one internal node per i */
        int d = (delta(k, n, i, i + 1) - delta(k, n, i, i - 1)) >= 0 ? 1 : -1;
        int dmin = delta(k, n, i, i - d);
        int lmax = 2;
        while (delta(k, n, i, i + lmax * d) > dmin) lmax <<= 1;
        int l = 0;
        for (int t = lmax >> 1; t >= 1; t >>= 1) if (delta(k, n, i, i + (l + t) * d) > dmin) l += t;
        int j = i + l * d;
        int dnode = delta(k, n, i, j);
        int s = 0, t = l;
        do { t = (t + 1) >> 1; if (delta(k, n, i, i + (s + t) * d) > dnode) s += t; } while (t > 1);
        int gamma = i + s * d + std::min(d, 0);
        int lo_i = std::min(i, j), hi_i = std::max(i, j);
        if (lo_i == gamma) { sc.children[2 * i] = ~gamma; sc.leaf_parent[gamma] = i; } else { sc.children[2 * i] = gamma; sc.parent[gamma] = i; }
        if (hi_i == gamma + 1) { sc.children[2 * i + 1] = ~(gamma + 1); sc.leaf_parent[gamma + 1] = i; } else { sc.children[2 * i + 1] = gamma + 1; sc.parent[gamma + 1] = i; }
    }
    /* refit: recursive post-order from the root (min/max are exact, order-free) */
    if (ni > 0) {
        std::vector<int> order; order.reserve(ni);
        std::vector<int> st; st.push_back(0);
        while (!st.empty()) { int nd = st.back(); st.pop_back(); order.push_back(nd); for (int c = 0; c < 2; ++c) { int ch = sc.children[2 * nd + c]; if (ch >= 0) st.push_back(ch); } }
        for (int idx = (int)order.size() - 1; idx >= 0; --idx) {
            int nd = order[idx];
            V3 mn[2], mx[2];
            for (int c = 0; c < 2; ++c) {
                int ch = sc.children[2 * nd + c];
                if (ch < 0) { uint32_t q = sc.sorted_prim[~ch]; mn[c] = sc.amin[q]; mx[c] = sc.amax[q]; }
                else { mn[c] = sc.node_min[ch]; mx[c] = sc.node_max[ch]; }
            }
            sc.node_min[nd] = mk(fminf(mn[0].x, mn[1].x), fminf(mn[0].y, mn[1].y), fminf(mn[0].z, mn[1].z));
            sc.node_max[nd] = mk(fmaxf(mx[0].x, mx[1].x), fmaxf(mx[0].y, mx[1].y), fmaxf(mx[0].z, mx[1].z));
        }
    }
    sc.built = true;
}

inline float safe_inv(float d) { float a = fabsf(d) < 1e-20f ? copysignf(1e-20f, d) : d; return 1.0f / a; }
/* Slab test of the two-child walk (prt_device.cuh:isect_box). pad is 0 except on axes the ray is parallel to (ray_pad below);
 * fmaf with a zero addend is the rounded product, so every other ray keeps the bits of (m - o) * inv. */
inline float isect_box(V3 o, V3 inv, V3 pad, V3 mn, V3 mx, float tmax) {
    float x0 = fmaf(mn.x - o.x, inv.x, -pad.x), x1 = fmaf(mx.x - o.x, inv.x, pad.x);
    float y0 = fmaf(mn.y - o.y, inv.y, -pad.y), y1 = fmaf(mx.y - o.y, inv.y, pad.y);
    float z0 = fmaf(mn.z - o.z, inv.z, -pad.z), z1 = fmaf(mx.z - o.z, inv.z, pad.z);
    float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax));
    tf = tf * 1.00000047683715820312f;
    return tn <= tf ? tn : -1.0f;
}
inline void child_box(const Scene& sc, int ch, V3& mn, V3& mx) {
    if (ch < 0) { uint32_t q = sc.sorted_prim[~ch]; mn = sc.amin[q]; mx = sc.amax[q]; }
    else { mn = sc.node_min[ch]; mx = sc.node_max[ch]; }
}

/* prt_device.cuh:make_raybox - a ray parallel to a slab (|d_k| < 1e-20, inv = +-1e20) that lies exactly in one of its planes
 * would "leave at t = 0"; on such axes the slab products are loosened by P = twice the largest t a point of the scene box
 * (the union of the root's two child boxes) can have on the ray's dominant axis */
inline V3 ray_pad(const Scene& sc, V3 o, V3 d, V3 inv) {
    V3 pad = mk(0.0f, 0.0f, 0.0f);
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    if ((ax < 1e-20f || ay < 1e-20f || az < 1e-20f) && sc.n_prims() > 1) {
        V3 lmn, lmx, rmn, rmx;
        child_box(sc, sc.children[0], lmn, lmx); child_box(sc, sc.children[1], rmn, rmx);
        const float ex = fmaxf(fabsf(fminf(lmn.x, rmn.x) - o.x), fabsf(fmaxf(lmx.x, rmx.x) - o.x));
        const float ey = fmaxf(fabsf(fminf(lmn.y, rmn.y) - o.y), fabsf(fmaxf(lmx.y, rmx.y) - o.y));
        const float ez = fmaxf(fabsf(fminf(lmn.z, rmn.z) - o.z), fabsf(fmaxf(lmx.z, rmx.z) - o.z));
        float P = (ax >= ay && ax >= az) ? ex * fabsf(inv.x) : (ay >= az ? ey * fabsf(inv.y) : ez * fabsf(inv.z));
        P = fminf(P * 2.0f, 1e30f);
        pad = mk(ax < 1e-20f ? copysignf(P, inv.x) : 0.0f, ay < 1e-20f ? copysignf(P, inv.y) : 0.0f, az < 1e-20f ? copysignf(P, inv.z) : 0.0f);
    }
    return pad;
}

/* deepest traversal stack any ray needed since the last reset, in entries of the DEVICE stack (its DONE sentinel included):
 * the device kernels hold PRT stack entries in shared + local memory (wavefront.cu: STACK_SMEM + STACK_LOCAL) */
std::atomic<int> g_max_stack{0};
inline void note_stack(int entries) { int cur = g_max_stack.load(std::memory_order_relaxed); while (entries > cur && !g_max_stack.compare_exchange_weak(cur, entries)) {} }

/* stack traversal, near child first; same visiting order as the device kernel so the
 * node / primitive counters are comparable */
Hit closest_bvh(const Scene& sc, V3 o, V3 d, uint64_t* n_nodes, uint64_t* n_tests) {
    Hit h; hit_init(h);
    int n = sc.n_prims();
    if (n == 0) return h;
    V3 inv = mk(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
    V3 pad = ray_pad(sc, o, d, inv);
    Shear rs = make_shear(d);
    V3 dn = normalize(d);
    int stack[512]; int sp = 0;
    int cur = n > 1 ? 0 : ~0;
    for (;;) {
        bool descend = false;
        if (cur >= 0) {
            if (n_nodes) ++*n_nodes;
            int cl = sc.children[2 * cur], cr = sc.children[2 * cur + 1];
            V3 lmn, lmx, rmn, rmx;
            child_box(sc, cl, lmn, lmx); child_box(sc, cr, rmn, rmx);
            float tl = isect_box(o, inv, pad, lmn, lmx, h.t);
            float tr = isect_box(o, inv, pad, rmn, rmx, h.t);
            if (tl >= 0.0f && tr >= 0.0f) {
                int nearc = cl, farc = cr;
                if (tr < tl) { nearc = cr; farc = cl; }
                stack[sp++] = farc; note_stack(sp + 1); cur = nearc; descend = true;
            } else if (tl >= 0.0f) { cur = cl; descend = true; }
            else if (tr >= 0.0f) { cur = cr; descend = true; }
        } else {
            if (n_tests) ++*n_tests;
            test_prim(sc, (int)sc.sorted_prim[~cur], o, d, dn, rs, h);
        }
        if (descend) continue;
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return h;
}

/* ---------------------------------------------------------------- 4-wide quantised nodes
 * Restatement of k_emit_wide (parallelraytracing_b200/csrc/lbvh.cu): every internal node at even depth
 * takes its children's children as 2..4 slots; child boxes are quantised to 8 bits on a per-node
 * power-of-two grid, floor/ceil with a 2^-6-step margin. The reference has no counterpart (no BVH at all). */
inline int dev_leaf_code(const Scene& sc, int pos) {
    uint32_t q = sc.sorted_prim[pos], ns = (uint32_t)sc.spheres.size(), nq = (uint32_t)sc.quads.size();
    uint32_t payload = q < ns ? q : q < ns + nq ? ((1u << 28) | (q - ns)) : ((2u << 28) | (uint32_t)pos);
    return ~(int)payload;
}
/* grid step of a wide node on one axis: the smallest power of two s with ext / s <= 253 (one step below the boxes is
 * kept free, see build_wide), at least 2 ulp of the larger bound (so that origin = mn - s is a different float) and >= 2^-60 */
inline float pow2_ceil_scale(float ext, float mn, float mx) {
    int e = -60;
    if (ext > 0.0f) { int e1; frexpf(ext * (1.0f / 253.0f), &e1); e = e1 > e ? e1 : e; }
    const float m = fmaxf(fabsf(mn), fabsf(mx));
    if (m > 0.0f) { int e2; frexpf(m, &e2); e2 -= 23; e = e2 > e ? e2 : e; }
    return ldexpf(1.0f, e);
}
void build_wide(Scene& sc) {
    int n = sc.n_prims(), ni = n > 1 ? n - 1 : 0;
    sc.wide.assign(ni, Scene::WNode()); sc.wide_valid.assign(ni, 0);
    for (int i = 0; i < ni; ++i) {
        /* greedy collapse by surface area, lbvh.cu:k_emit_wide: start from the two children, twice open the internal slot
         * with the largest box (ties: lowest slot), its two children take its place */
        V3 lo[4], hi[4]; int code[4]; int raw[4]; int ns = 2;
        for (int k = 0; k < 2; ++k) {
            raw[k] = sc.children[2 * i + k];
            child_box(sc, raw[k], lo[k], hi[k]);
        }
        for (int round = 0; round < 2; ++round) {
            int s = -1; float best = -1.0f;
            for (int c = 0; c < ns; ++c)
                if (raw[c] >= 0) {
                    const float dx = hi[c].x - lo[c].x, dy = hi[c].y - lo[c].y, dz = hi[c].z - lo[c].z;
                    const float ar = (dx * dy + dy * dz) + dz * dx;
                    if (ar > best) { best = ar; s = c; }
                }
            if (s < 0) break;
            for (int c = ns; c > s + 1; --c) { raw[c] = raw[c - 1]; lo[c] = lo[c - 1]; hi[c] = hi[c - 1]; }
            const int node = raw[s];
            raw[s] = sc.children[2 * node]; raw[s + 1] = sc.children[2 * node + 1];
            child_box(sc, raw[s], lo[s], hi[s]); child_box(sc, raw[s + 1], lo[s + 1], hi[s + 1]);
            ++ns;
        }
        for (int c = 0; c < ns; ++c) code[c] = raw[c] < 0 ? dev_leaf_code(sc, ~raw[c]) : raw[c];
        Scene::WNode& w = sc.wide[i];
        for (int k = 0; k < 3; ++k) {
            float mn = comp(lo[0], k), mx = comp(hi[0], k);
            for (int c = 1; c < ns; ++c) { mn = fminf(mn, comp(lo[c], k)); mx = fmaxf(mx, comp(hi[c], k)); }
            float sc_k = pow2_ceil_scale(mx - mn, mn, mx);
            /* The grid starts ONE STEP BELOW the lowest box: the floor of a box that starts at the node's own minimum then
             * lands on step 0 with a margin of >= half a step instead of being clamped there with none (the fused decode
             * in the walk rounds by up to 2^-9 step; a low face without margin lost corner-grazing hits). */
            const float org = mn - sc_k;
            uint32_t ql = 0, qh = 0;
            for (int c = 0; c < 4; ++c) {
                uint32_t a = 255u, b = 0u;
                if (c < ns) {
                    float fl = floorf((comp(lo[c], k) - org) / sc_k - 0.015625f), fh = ceilf((comp(hi[c], k) - org) / sc_k + 0.015625f);
                    a = (uint32_t)fminf(fmaxf(fl, 0.0f), 255.0f); b = (uint32_t)fminf(fmaxf(fh, 0.0f), 255.0f);
                }
                ql |= a << (8 * c); qh |= b << (8 * c);
            }
            w.org[k] = org; w.scale[k] = sc_k * 32768.0f; w.qlo[k] = ql; w.qhi[k] = qh;
        }
        for (int c = 0; c < 4; ++c) w.code[c] = c < ns ? code[c] : code[0];
        sc.wide_valid[i] = 1;
    }
}

inline float bits_f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
inline uint32_t f_bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

/* Restatement of k_extend_wide's walk (wavefront.cu): same decode arithmetic (explicit fmaf, same order), nearest
 * hit child next, the other hit children pushed in slot order; leaf codes are the device's. Counters comparable. */
Hit closest_wide(const Scene& sc, V3 o, V3 d, uint64_t* n_nodes, uint64_t* n_tests) {
    Hit h; hit_init(h);
    int n = sc.n_prims();
    if (n == 0) return h;
    if (n == 1) return closest_bvh(sc, o, d, n_nodes, n_tests);
    V3 inv = mk(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
    Shear rs = make_shear(d);
    V3 dn = normalize(d);
    const int ns = (int)sc.spheres.size();
    std::vector<int> stack; stack.reserve(128);
    const int DONE = (int)0x80000000;
    int cur = 0;
    const bool nx = inv.x < 0.0f, ny = inv.y < 0.0f, nz = inv.z < 0.0f;
    const float e = 2.384185791015625e-07f;
    for (;;) {
        if (cur >= 0) {
            if (n_nodes) ++*n_nodes;
            const Scene::WNode& w = sc.wide[cur];
            const float Bx = w.scale[0] * inv.x, By = w.scale[1] * inv.y, Bz = w.scale[2] * inv.z;
            const float Ax = (w.org[0] - o.x) * inv.x, Ay = (w.org[1] - o.y) * inv.y, Az = (w.org[2] - o.z) * inv.z;
            const float nAx = fmaf(-fabsf(Ax), e, Ax) - Bx, fAx = fmaf(fabsf(Ax), e, Ax) - Bx;
            const float nAy = fmaf(-fabsf(Ay), e, Ay) - By, fAy = fmaf(fabsf(Ay), e, Ay) - By;
            const float nAz = fmaf(-fabsf(Az), e, Az) - Bz, fAz = fmaf(fabsf(Az), e, Az) - Bz;
            const uint32_t wnx = nx ? w.qhi[0] : w.qlo[0], wfx = nx ? w.qlo[0] : w.qhi[0];
            const uint32_t wny = ny ? w.qhi[1] : w.qlo[1], wfy = ny ? w.qlo[1] : w.qhi[1];
            const uint32_t wnz = nz ? w.qhi[2] : w.qlo[2], wfz = nz ? w.qlo[2] : w.qhi[2];
            uint32_t key[4]; bool hit[4];
            for (int k = 0; k < 4; ++k) {
                auto g = [&](uint32_t word) { return bits_f(0x3f800000u | (((word >> (8 * k)) & 255u) << 8)); };
                float tn = fmaxf(fmaxf(fmaf(g(wnx), Bx, nAx), fmaf(g(wny), By, nAy)), fmaxf(fmaf(g(wnz), Bz, nAz), 0.0f));
                float tf = fminf(fminf(fmaf(g(wfx), Bx, fAx), fmaf(g(wfy), By, fAy)), fminf(fmaf(g(wfz), Bz, fAz), h.t));
                tf = tf * 1.00000047683715820312f;
                hit[k] = tn <= tf;
                key[k] = hit[k] ? ((f_bits(tn) & ~3u) | (uint32_t)k) : 0xffffffffu;
            }
            uint32_t kmin = std::min(std::min(key[0], key[1]), std::min(key[2], key[3]));
            if (kmin != 0xffffffffu) {
                for (int k = 0; k < 4; ++k) if (hit[k] && key[k] > kmin) stack.push_back(w.code[k]);
                note_stack((int)stack.size() + 1);
                cur = w.code[kmin & 3u];
                continue;
            }
        } else {
            if (n_tests) ++*n_tests;
            uint32_t payload = (uint32_t)~cur, type = payload >> 28, idx = payload & 0x0fffffffu;
            int q = type == 0 ? (int)idx : type == 1 ? ns + (int)idx : (int)sc.sorted_prim[idx];
            test_prim(sc, q, o, d, dn, rs, h);
        }
        if (stack.empty()) break;
        cur = stack.back(); stack.pop_back();
        if (cur == DONE) break;
    }
    return h;
}

/* ---------------------------------------------------------------- surface + materials */
struct Surface { V3 p, n; bool front; int mat; };

Surface make_surface(const Scene& sc, V3 o, V3 d, const Hit& h) {
    Surface s;
    int ns = (int)sc.spheres.size(), nq = (int)sc.quads.size();
    if (h.q < ns) {
        const prt_sphere& sp = sc.spheres[h.q];
        V3 c = mk(sp.cx, sp.cy, sp.cz);
        V3 pl = add(sub(o, c), mul(normalize(d), h.t));        /* local hit point (shape.h:194) */
        V3 n = normalize(pl);
        s.front = h.u != 0.0f;
        if (!s.front) n = neg(n);
        s.n = normalize(n);                                     /* TransformNormal, primitive.cpp:39 */
        s.p = add(pl, c);                                       /* Mat * p', primitive.cpp:38 */
        s.mat = sp.material;
    } else if (h.q < ns + nq) {
        const prt_quad& q = sc.quads[h.q - ns];
        s.p = add(o, mul(d, h.t));
        s.front = h.u != 0.0f;
        V3 n = mk(q.nx, q.ny, q.nz);
        s.n = s.front ? n : neg(n);
        s.mat = q.material;
    } else {
        const Tri& t = sc.tris[h.q - ns - nq];
        float b0 = (1.0f - h.u) - h.v;
        s.p = add(add(mul(t.v[0], b0), mul(t.v[1], h.u)), mul(t.v[2], h.v));
        V3 n = add(add(mul(t.n[0], b0), mul(t.n[1], h.u)), mul(t.n[2], h.v));
        if (dot(n, d) > 0.0f) { n = neg(n); s.front = false; } else s.front = true;
        s.n = normalize(n);
        s.mat = t.mat;
    }
    return s;
}

inline V3 reflect(V3 i, V3 n) { float k = dot(n, i); return sub(i, mul(mul(n, k), 2.0f)); }

/* MaterialHandle::{Emit,Scatter} (material.h:139-161); out_d already through Ray::Normalize
 * (cpu/renderer.cpp:84). */
bool scatter(const Scene& sc, int mat, V3 d, const Surface& s, Rng& rng, V3& atten, V3& out_d, V3& emit) {
    emit = mk(0.f, 0.f, 0.f);
    if (mat < 0 || mat >= (int)sc.mats.size()) return false;
    const prt_material& m = sc.mats[mat];
    V3 dir;
    if (m.type == PRT_MAT_LAMBERTIAN) {                          /* material.h:16-31 */
        V3 ruv = rng_unit_vector(rng);
        dir = add(s.n, ruv);
        if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = s.n;
        dir = normalize(dir);
        atten = mk(m.r, m.g, m.b);
    } else if (m.type == PRT_MAT_METAL) {                        /* material.h:48-57 */
        V3 r = normalize(reflect(d, s.n));
        V3 ruv = rng_unit_vector(rng);
        r = add(r, mul(ruv, m.param));
        dir = normalize(r);
        atten = mk(m.r, m.g, m.b);
        if (!(dot(dir, s.n) > 0.0f)) return false;
    } else if (m.type == PRT_MAT_DIELECTRIC) {                   /* material.h:76-95 */
        atten = mk(1.f, 1.f, 1.f);
        float ri = s.front ? (1.0f / m.param) : m.param;
        float cos_t = fminf(dot(neg(d), s.n), 1.0f);
        float sin_t = sqrtf(1.0f - cos_t * cos_t);
        bool cannot = ri * sin_t > 1.0f;
        float r0 = (1.0f - ri) / (1.0f + ri);
        r0 = r0 * r0;
        float x = 1.0f - cos_t;
        float x5 = ((x * x) * (x * x)) * x;
        float fres = r0 + (1.0f - r0) * x5;
        bool refl = cannot;
        if (!refl) refl = fres > rng_uniform(rng);
        if (refl) dir = reflect(d, s.n);
        else {                                                   /* math.h:45-50 */
            float c2 = fminf(dot(neg(d), s.n), 1.0f);
            V3 perp = mul(add(d, mul(s.n, c2)), ri);
            float k = -sqrtf(fabsf(1.0f - dot(perp, perp)));
            dir = add(perp, mul(s.n, k));
        }
    } else {
        if (m.type == PRT_MAT_EMISSIVE) emit = mk(m.r, m.g, m.b); /* material.h:119-127 */
        return false;
    }
    out_d = normalize(dir);
    return true;
}

struct Cam { V3 pos, front, right, up; float tan_half, aspect, w, h; int W, H; };
Cam make_cam(const prt_camera& k) {
    Cam c;
    c.pos = mk(k.pos[0], k.pos[1], k.pos[2]); c.front = mk(k.front[0], k.front[1], k.front[2]);
    c.right = mk(k.right[0], k.right[1], k.right[2]); c.up = mk(k.up[0], k.up[1], k.up[2]);
    c.tan_half = k.tan_half_fov; c.w = (float)k.width; c.h = (float)k.height; c.aspect = c.w / c.h;
    c.W = k.width; c.H = k.height;
    return c;
}
/* Camera::GetCameraRay, camera.h:104-132 */
V3 camera_dir(const Cam& c, float px, float py) {
    float ndc_x = (px / c.w) * 2.0f - 1.0f;
    float ndc_y = 1.0f - (py / c.h) * 2.0f;
    V3 dc = normalize(mk(ndc_x * c.aspect * c.tan_half, ndc_y * c.tan_half, -1.0f));
    V3 w = add(add(mul(c.right, dc.x), mul(c.up, dc.y)), mul(neg(c.front), dc.z));
    return normalize(w);
}

/* CPURenderer::TraceRay (cpu/renderer.cpp:59-103), iterative: at most max_depth closest-hit queries */
V3 trace_path(const Scene& sc, V3 o, V3 d, uint32_t key, int max_depth, bool use_bvh, uint64_t* rays) {
    V3 beta = mk(1.f, 1.f, 1.f), L = mk(0.f, 0.f, 0.f);
    auto clampL = [&](V3 v) { return sc.clamp < INFINITY ? mk(fminf(v.x, sc.clamp), fminf(v.y, sc.clamp), fminf(v.z, sc.clamp)) : v; };
    for (int bounce = 0; bounce < max_depth; ++bounce) {
        ++*rays;
        Hit h = use_bvh ? closest_bvh(sc, o, d, nullptr, nullptr) : closest_brute(sc, o, d, nullptr);
        if (h.q < 0) { L = add(L, clampL(mulv(beta, mk(sc.sky[0], sc.sky[1], sc.sky[2])))); break; }   /* renderer.cpp:99 */
        Surface s = make_surface(sc, o, d, h);
        Rng rng = rng_for_bounce(key, (uint32_t)bounce);
        V3 atten, nd, emit;
        bool ok = scatter(sc, s.mat, d, s, rng, atten, nd, emit);
        L = add(L, clampL(mulv(beta, emit)));                                                   /* renderer.cpp:74 */
        if (!ok) break;
        beta = mulv(beta, atten);                                                               /* renderer.cpp:93 */
        /* options of the device path, off by default (the reference has neither): Russian roulette */
        if (sc.rr_start >= 0 && bounce >= sc.rr_start && bounce + 1 < max_depth) {
            const float p = fmaxf(fminf(fmaxf(fmaxf(beta.x, beta.y), beta.z), 1.0f), 0.05f);
            if (rng_uniform(rng) > p) break;
            beta = mk(beta.x / p, beta.y / p, beta.z / p);
        }
        o = s.p; d = nd;
    }
    return L;
}

}  // namespace

/* ==================================================================== C ABI for ctypes */
extern "C" {

void* po_scene_create(const prt_sphere* spheres, int ns, const prt_quad* quads, int nq,
                      const float* tri_verts, const float* tri_normals, const int32_t* tri_material,
                      int tri_material_default, int nt, int tri_prim_base,
                      const prt_material* mats, int nm, const float* sky) {
    Scene* sc = new Scene();
    sc->spheres.assign(spheres, spheres + ns);
    sc->quads.assign(quads, quads + nq);
    sc->mats.assign(mats, mats + nm);
    sc->tris.resize(nt);
    for (int64_t t = 0; t < nt; ++t) {
        Tri& T = sc->tris[t];
        for (int k = 0; k < 3; ++k) {
            T.v[k] = mk(tri_verts[9 * t + 3 * k], tri_verts[9 * t + 3 * k + 1], tri_verts[9 * t + 3 * k + 2]);
            T.n[k] = mk(tri_normals[9 * t + 3 * k], tri_normals[9 * t + 3 * k + 1], tri_normals[9 * t + 3 * k + 2]);
        }
        T.mat = tri_material ? tri_material[t] : tri_material_default;
    }
    sc->tri_prim_base = tri_prim_base;
    if (sky) { sc->sky[0] = sky[0]; sc->sky[1] = sky[1]; sc->sky[2] = sky[2]; }
    return sc;
}
void po_scene_destroy(void* s) { delete (Scene*)s; }
void po_set_estimator(void* s, int rr_start, float clamp) { Scene& sc = *(Scene*)s; sc.rr_start = rr_start; sc.clamp = clamp > 0.f ? clamp : INFINITY; }
int po_num_prims(void* s) { return ((Scene*)s)->n_prims(); }

void po_set_builder(void* s, int builder, int radius) { Scene& sc = *(Scene*)s; sc.builder = builder; sc.ploc_radius = radius; }
int po_ploc_rounds(void* s) { return ((Scene*)s)->ploc_rounds; }
int po_bvh_height(void* s) { return ((Scene*)s)->height; }
int po_builder_used(void* s) { return ((Scene*)s)->builder_used; }
void po_reset_max_stack(void) { g_max_stack.store(0); }
int po_get_max_stack(void) { return g_max_stack.load(); }
void po_build_lbvh(void* s) { build_lbvh(*(Scene*)s); build_wide(*(Scene*)s); }

/* 4-wide nodes in the device layout: 16 words per internal node {ox,oy,oz,sx, code[4], qlo[3], qhi[3], sy,sz}; valid[i] = even depth */
void po_get_wide(void* s, uint32_t* words, uint8_t* valid) {
    Scene& sc = *(Scene*)s;
    for (size_t i = 0; i < sc.wide.size(); ++i) {
        const Scene::WNode& w = sc.wide[i]; uint32_t* o = words + 16 * i;
        if (valid) valid[i] = sc.wide_valid[i];
        if (!words) continue;
        o[0] = f_bits(w.org[0]); o[1] = f_bits(w.org[1]); o[2] = f_bits(w.org[2]); o[3] = f_bits(w.scale[0]);
        for (int c = 0; c < 4; ++c) o[4 + c] = (uint32_t)w.code[c];
        for (int k = 0; k < 3; ++k) { o[8 + k] = w.qlo[k]; o[11 + k] = w.qhi[k]; }
        o[14] = f_bits(w.scale[1]); o[15] = f_bits(w.scale[2]);
    }
}

/* same outputs as prt_debug_lbvh */
void po_get_lbvh(void* s, uint32_t* morton, uint32_t* sorted_prim, int32_t* node_children, int32_t* node_parent,
                 float* node_bounds, float* scene_bounds) {
    Scene& sc = *(Scene*)s;
    int n = sc.n_prims(), ni = n > 1 ? n - 1 : 0;
    if (scene_bounds) memcpy(scene_bounds, sc.bounds, sizeof sc.bounds);
    if (morton) memcpy(morton, sc.morton.data(), (size_t)n * 4);
    if (sorted_prim) memcpy(sorted_prim, sc.sorted_prim.data(), (size_t)n * 4);
    if (node_children) memcpy(node_children, sc.children.data(), (size_t)ni * 8);
    if (node_parent) memcpy(node_parent, sc.parent.data(), (size_t)ni * 4);
    if (node_bounds) for (int i = 0; i < ni; ++i) {
        V3 mn, mx; float* o = node_bounds + 12 * (size_t)i;
        child_box(sc, sc.children[2 * i], mn, mx);
        o[0] = mn.x; o[1] = mn.y; o[2] = mn.z; o[3] = mx.x; o[4] = mx.y; o[5] = mx.z;
        child_box(sc, sc.children[2 * i + 1], mn, mx);
        o[6] = mn.x; o[7] = mn.y; o[8] = mn.z; o[9] = mx.x; o[10] = mx.y; o[11] = mx.z;
    }
}

/* closest hit for a ray batch. use_bvh: 0 brute force, 1 LBVH traversal, 2 the 4-wide quantised nodes. counters[0..1] += nodes, prim tests */
void po_trace_batch(void* s, const prt_ray* rays, int n, prt_hit* out, int use_bvh, uint64_t* counters, int n_threads) {
    Scene& sc = *(Scene*)s;
    if (n_threads < 1) n_threads = 1;
    std::vector<uint64_t> nn(n_threads, 0), np(n_threads, 0);
    auto work = [&](int tid) {
        for (int i = tid; i < n; i += n_threads) {
            V3 o = mk(rays[i].ox, rays[i].oy, rays[i].oz), d = mk(rays[i].dx, rays[i].dy, rays[i].dz);
            Hit h = use_bvh == 2 ? closest_wide(sc, o, d, &nn[tid], &np[tid])
                  : use_bvh ? closest_bvh(sc, o, d, &nn[tid], &np[tid]) : closest_brute(sc, o, d, &np[tid]);
            prt_hit r; memset(&r, 0, sizeof r);
            if (h.q < 0) { r.t = h.t; r.prim_id = -1; r.material = -1; }
            else {
                Surface sf = make_surface(sc, o, d, h);
                bool is_tri = h.q >= (int)(sc.spheres.size() + sc.quads.size());
                r.t = h.t; r.prim_id = h.prim_id; r.b1 = is_tri ? h.u : 0.f; r.b2 = is_tri ? h.v : 0.f;
                r.px = sf.p.x; r.py = sf.p.y; r.pz = sf.p.z; r.nx = sf.n.x; r.ny = sf.n.y; r.nz = sf.n.z;
                r.front = sf.front ? 1 : 0; r.material = sf.mat;
            }
            out[i] = r;
        }
    };
    if (n_threads == 1) work(0);
    else { std::vector<std::thread> th; for (int t = 0; t < n_threads; ++t) th.emplace_back(work, t); for (auto& t : th) t.join(); }
    if (counters) for (int t = 0; t < n_threads; ++t) { counters[0] += nn[t]; counters[1] += np[t]; }
}

/* camera directions for all pixels (W*H*3 floats), pixel centres */
void po_camera_dirs(const prt_camera* cam, float* out) {
    Cam c = make_cam(*cam);
    for (int y = 0; y < c.H; ++y) for (int x = 0; x < c.W; ++x) {
        V3 d = camera_dir(c, (float)x + 0.5f, (float)y + 0.5f);
        float* o = out + 3 * ((size_t)y * c.W + x); o[0] = d.x; o[1] = d.y; o[2] = d.z;
    }
}

/* adds samples [sample_begin, sample_begin+spp) of every pixel into accum (W*H*3), in sample
 * order per pixel (Film::AddSample). Rows are split across n_threads. Returns #closest-hit queries. */
uint64_t po_render(void* s, const prt_camera* cam, int spp, int max_depth, uint64_t seed, int sample_begin,
                   float* accum, int use_bvh, int n_threads, int y_begin, int y_end) {
    Scene& sc = *(Scene*)s;
    Cam c = make_cam(*cam);
    if (n_threads < 1) n_threads = 1;
    if (y_end <= 0 || y_end > c.H) y_end = c.H;
    if (y_begin < 0) y_begin = 0;
    std::vector<uint64_t> rays(n_threads, 0);
    std::atomic<int> next_row(y_begin);
    auto work = [&](int tid) {
        for (;;) {
            int y = next_row.fetch_add(1);
            if (y >= y_end) break;
            for (int x = 0; x < c.W; ++x) {
                uint32_t pixel = (uint32_t)(y * c.W + x);
                V3 d0 = camera_dir(c, (float)x + 0.5f, (float)y + 0.5f);
                float* a = accum + 3 * (size_t)pixel;
                for (int sidx = 0; sidx < spp; ++sidx) {
                    uint32_t key = path_key(seed, pixel, (uint32_t)(sample_begin + sidx));
                    V3 L = trace_path(sc, c.pos, d0, key, max_depth, use_bvh != 0, &rays[tid]);
                    a[0] += L.x; a[1] += L.y; a[2] += L.z;
                }
            }
        }
    };
    if (n_threads == 1) work(0);
    else { std::vector<std::thread> th; for (int t = 0; t < n_threads; ++t) th.emplace_back(work, t); for (auto& t : th) t.join(); }
    uint64_t total = 0; for (uint64_t r : rays) total += r;
    return total;
}

/* one path (debug / unit parity): radiance of sample `sample` of pixel (x,y) */
void po_trace_pixel(void* s, const prt_camera* cam, int x, int y, int sample, int max_depth, uint64_t seed, int use_bvh, float* out_rgb) {
    Scene& sc = *(Scene*)s; Cam c = make_cam(*cam);
    uint64_t rays = 0;
    V3 d0 = camera_dir(c, (float)x + 0.5f, (float)y + 0.5f);
    V3 L = trace_path(sc, c.pos, d0, path_key(seed, (uint32_t)(y * c.W + x), (uint32_t)sample), max_depth, use_bvh != 0, &rays);
    out_rgb[0] = L.x; out_rgb[1] = L.y; out_rgb[2] = L.z;
}

/* MaterialHandle::{Emit,Scatter} known answers (material.h:139-161), one record: the hit normal already faces the ray,
 * the RNG is the (seed, pixel, sample, bounce) stream; out = {ok, atten.rgb, out_dir.xyz (0 when the path ends), emit.rgb}.
 * Same record layout as oracle/ref_harness.cpp:ref_scatter and prt_debug_scatter. */
void po_scatter(const prt_material* m, const float* in_dir, const float* normal, int front, uint64_t seed, uint32_t pixel,
                uint32_t sample, uint32_t bounce, float* out10) {
    Scene sc; sc.mats.assign(m, m + 1);
    Surface sf; sf.p = mk(0.f, 0.f, 0.f); sf.n = mk(normal[0], normal[1], normal[2]); sf.front = front != 0; sf.mat = 0;
    Rng rng = rng_for_bounce(path_key(seed, pixel, sample), bounce);
    V3 atten = mk(0.f, 0.f, 0.f), od = mk(0.f, 0.f, 0.f), emit;
    bool ok = scatter(sc, 0, mk(in_dir[0], in_dir[1], in_dir[2]), sf, rng, atten, od, emit);
    if (!ok) od = mk(0.f, 0.f, 0.f);
    out10[0] = ok ? 1.f : 0.f; out10[1] = atten.x; out10[2] = atten.y; out10[3] = atten.z;
    out10[4] = od.x; out10[5] = od.y; out10[6] = od.z; out10[7] = emit.x; out10[8] = emit.y; out10[9] = emit.z;
}

/* the RNG stream of (seed, pixel, sample, bounce): first n rand() values (31 bit) */
void po_rng_stream(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, int n, uint32_t* out) {
    Rng r = rng_for_bounce(path_key(seed, pixel, sample), bounce);
    for (int i = 0; i < n; ++i) out[i] = rng_next31(r);
}

/* Film::UpdateDisplay (film.cu:134-194) with a uniform per-pixel weight `spp` */
void po_resolve_display(const float* accum, int npix, float spp, float exposure, float gamma, uint8_t* out) {
    float inv_gamma = 1.0f / gamma;
    for (int i = 0; i < npix; ++i) {
        float rgb[3] = {0.f, 0.f, 0.f};
        if (spp > 0.0f) {
            float inv_w = 1.0f / spp;
            for (int c = 0; c < 3; ++c) {
                float v = accum[3 * (size_t)i + c] * inv_w;
                float x = v * exposure; x = x / (1.0f + x);
                rgb[c] = powf(x, inv_gamma);
            }
        }
        for (int c = 0; c < 3; ++c) {
            float v = fmaxf(0.0f, fminf(1.0f, rgb[c]));
            out[4 * (size_t)i + c] = (uint8_t)(v * 255.0f + 0.5f);
        }
        out[4 * (size_t)i + 3] = 255;
    }
}

}  /* extern "C" */
