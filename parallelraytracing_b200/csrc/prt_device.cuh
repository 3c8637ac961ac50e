// prt_device.cuh — device-side math of the hot path: vectors, counter RNG, the three
// primitive tests, the four materials, the camera ray.
//
// Floating-point contract (DESIGN.md "Numerics"): this translation unit is compiled with
// -fmad=false, so `a*b+c` is NEVER contracted; every fused multiply-add below is an explicit
// fmaf(). The CPU oracle (oracle/prt_oracle.cpp) restates the same formulas with the same
// operation order (-ffp-contract=off), IEEE sqrt and divide on both sides, so integer
// results (hit ids, Morton codes) are bit-exact and floats agree to the last bit in practice.
//
// Reference semantics mirrored (paths relative to /root/reference):
//   camera ray        src/core/camera.h:104-132
//   sphere ("Circle") src/core/shape.h:157-203
//   quad              src/core/shape.h:213-239
//   triangle          src/core/shape.h:262-303 (Moller-Trumbore there; watertight here)
//   closest hit rule  src/core/primitive.cpp:21-59 (strict <, lowest list index wins ties)
//   materials         src/core/material.h:11-161, src/core/math.h:10-50
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define PRT_DEV __device__ __forceinline__

// Checked build (-DPRT_CHECKED=1, libprt_b200_checked.so): every index that device code derives from data - traversal stack
// pointers, queue slots, node / leaf codes, PLOC output positions and node ids, radix-sort scatter positions - is tested
// before the access; a violation sets a bit in a device-global word (one per translation unit, read by prt_debug_check_flags)
// and the access is skipped. compute-sanitizer is closed on this GPU pool, so this is the repo's own memcheck: the tests run
// the small configurations of every kernel family through the checked library and require the word to stay zero.
#define PRT_CHK_STACK   1u   /* traversal stack pointer beyond STACK_TOTAL                   */
#define PRT_CHK_QUEUE   2u   /* ray / hit queue slot beyond the wave's capacity             */
#define PRT_CHK_NODE    4u   /* node index or leaf code outside the BVH / primitive arrays  */
#define PRT_CHK_PLOC    8u   /* PLOC: output position, node id or shared-memory slot        */
#define PRT_CHK_SORT   16u   /* radix sort: scatter position beyond n                        */
#define PRT_CHK_REFIT  32u   /* refit / wide emission: node index beyond n - 1               */
#if PRT_CHECKED
static __device__ unsigned int g_prt_check_flags = 0u;
#define PRT_CHECK(cond, bit) ((cond) ? true : (atomicOr(&g_prt_check_flags, (bit)), false))
#else
#define PRT_CHECK(cond, bit) (true)
#endif

namespace prt {

// ------------------------------------------------------------------ vectors
struct V3 { float x, y, z; };

PRT_DEV V3 mk(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
PRT_DEV V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
PRT_DEV V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
PRT_DEV V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
PRT_DEV V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
PRT_DEV V3 neg(V3 a) { return mk(-a.x, -a.y, -a.z); }
// glm::dot(vec3): (x*x' + y*y') + z*z', unfused.
PRT_DEV float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
PRT_DEV V3 cross(V3 a, V3 b) {
    return mk(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
// glm::normalize: v * inversesqrt(dot(v,v)), inversesqrt = 1/sqrt (IEEE both).
PRT_DEV V3 normalize(V3 v) { float s = 1.0f / sqrtf(dot(v, v)); return v * s; }
PRT_DEV V3 xyz(float4 f) { return mk(f.x, f.y, f.z); }

// ------------------------------------------------------------------ tolerance-mode float math (option fast_math)
// The default (FAST = false) is the bit-exact contract above: IEEE divide and square root in the oracle's operation order.
// FAST = true replaces them ON THE SHADING SIDE ONLY - camera ray, surface normal, the four materials, the rejection-sampled
// unit vector - by the SFU approximations (MUFU.RCP / RSQ / SQRT, <= 2 ulp, one instruction instead of a 6-10 instruction
// refinement sequence). Intersection and traversal stay exact, so closest-hit ids, t and every integer output are unchanged;
// radiance moves within the tolerance the north star states (per-pixel RMSE <= 1e-3, max relative error <= 1e-2; measured and
// asserted in tests/test_gpu_parity.py::test_fast_math_mode).
PRT_DEV float fast_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
PRT_DEV float fast_rsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
PRT_DEV float fast_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
template <bool FAST> PRT_DEV float div_t(float a, float b) { if (FAST) return a * fast_rcp(b); return a / b; }
template <bool FAST> PRT_DEV float sqrt_t(float x) { if (FAST) return fast_sqrt(x); return sqrtf(x); }
template <bool FAST> PRT_DEV V3 normalize_t(V3 v) { if (FAST) return v * fast_rsqrt(dot(v, v)); return normalize(v); }

// ------------------------------------------------------------------ exact division by a launch constant
// floor(n / d) for n < 2^31 as one 32x32->64 multiply and a shift (Granlund & Montgomery 1994,
// m = ceil(2^(31+L) / d), L = ceil(log2 d)); replaces the ~25-instruction integer division.
struct FastDiv { uint32_t mul, shift, d; };
static inline FastDiv make_fastdiv(uint32_t d) {
    FastDiv f; f.d = d;
    uint32_t L = 0; while ((1ull << L) < d) ++L;
    f.shift = 31u + L;
    f.mul = (uint32_t)(((1ull << f.shift) + d - 1) / d);
    return f;
}
PRT_DEV uint32_t fdiv(uint32_t n, const FastDiv& f) { return (uint32_t)(((unsigned long long)n * f.mul) >> f.shift); }

// ------------------------------------------------------------------ RNG
// Counter-free PCG-RXS-M-XS-32 stream per (path, bounce). Random() mirrors the host
// branch of the reference: std::rand() / (RAND_MAX + 1.0f) with a 31-bit rand()
// (src/core/math.h:15); the int->float conversion rounds to nearest, so 1.0f can occur.
PRT_DEV uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x21f0aaadu; x ^= x >> 15; x *= 0x735a2d97u; x ^= x >> 15;
    return x;
}
PRT_DEV uint32_t path_key(uint32_t seed_lo, uint32_t seed_hi, uint32_t pixel, uint32_t sample) {
    uint32_t k = hash32(seed_hi ^ 0x9e3779b9u);
    k = hash32(seed_lo ^ k);
    k = hash32(sample ^ k);
    k = hash32(pixel ^ k);
    return k;
}
struct Rng { uint32_t s; };
PRT_DEV Rng rng_for_bounce(uint32_t key, uint32_t bounce) {
    Rng r; r.s = hash32(key + 0x9e3779b9u * (bounce + 1u)); return r;
}
PRT_DEV uint32_t rng_next31(Rng& r) {
    r.s = r.s * 747796405u + 2891336453u;
    uint32_t w = ((r.s >> ((r.s >> 28) + 4u)) ^ r.s) * 277803737u;
    w = (w >> 22) ^ w;
    return w >> 1;
}
PRT_DEV float rng_uniform(Rng& r) {
    return __int2float_rn((int)rng_next31(r)) * 4.656612873077392578125e-10f;  // / 2^31
}
// Random(-1, 1): min + (max - min) * Random()   (math.h:20-23)
// One FFMA: x*2^-31, the doubling and x*2^-30 are all exact (power-of-two scalings of a 24-bit significand), so
// fmaf(x, 2^-30, -1) rounds the same real number once, exactly like the three-operation form does.
PRT_DEV float rng_pm1(Rng& r) {
    return fmaf(__int2float_rn((int)rng_next31(r)), 9.31322574615478515625e-10f, -1.0f);
}
// RandomUnitVector: rejection in the cube (math.h:26-36)
template <bool FAST = false>
PRT_DEV V3 rng_unit_vector(Rng& r) {
    for (;;) {
        float x = rng_pm1(r);
        float y = rng_pm1(r);
        float z = rng_pm1(r);
        V3 p = mk(x, y, z);
        float l2 = dot(p, p);
        if (1e-8f < l2 && l2 <= 1.0f) {
            if (FAST) return p * fast_rsqrt(l2);
            float l = sqrtf(l2);
            return mk(p.x / l, p.y / l, p.z / l);
        }
    }
}

// 256-bit read-only load (sm_100 LDG.E.256): a divergent node fetch costs one L1 wavefront per 128-byte
// line per INSTRUCTION, so 2 x 32 B instead of 4 x 16 B halves the L1tex work of a 64-byte BVH node.
PRT_DEV void ldg8(const void* p, float4& a, float4& b) {
    unsigned r0, r1, r2, r3, r4, r5, r6, r7;
    asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7) : "l"(p));
    a = make_float4(__uint_as_float(r0), __uint_as_float(r1), __uint_as_float(r2), __uint_as_float(r3));
    b = make_float4(__uint_as_float(r4), __uint_as_float(r5), __uint_as_float(r6), __uint_as_float(r7));
}
// 256-bit store (sm_100 STG.E.256); p must be 32-byte aligned
PRT_DEV void stg8(void* p, float4 a, float4 b) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p),
                 "r"(__float_as_uint(a.x)), "r"(__float_as_uint(a.y)), "r"(__float_as_uint(a.z)), "r"(__float_as_uint(a.w)),
                 "r"(__float_as_uint(b.x)), "r"(__float_as_uint(b.y)), "r"(__float_as_uint(b.z)), "r"(__float_as_uint(b.w)) : "memory");
}

// ------------------------------------------------------------------ scene records
// Device layouts (16-byte planes, fetched with 128-bit loads).
struct SphereRec { float4 a; float4 b; };          // a={c.xyz, r}  b={tmin, mat, prim_id, 0}
struct QuadRec   { float4 a, b, c, d; };            // a={c.xyz,hw} b={u.xyz,hh} c={v.xyz,tmin} d={n.xyz, bits(prim_id)}
struct MatRec    { float4 a; float4 b; };           // a={r,g,b,param} b={bits(type),0,0,0}
struct BvhNode   { float4 lmin, lmax, rmin, rmax; };// .w of lmin/lmax = bits(left/right child code)

// 4-wide node with child boxes quantised to 8 bits on a per-node grid (64 bytes, two 256-bit loads for FOUR
// children; the two-child node above costs the same two loads for two). Child c's box on axis k is
//   [origin_k + qlo_k[c]*s_k , origin_k + qhi_k[c]*s_k],  s_k = scale_k * 2^-15 (a power of two),
// a superset of the true box by at least 2^-6 grid steps (the builder floors/ceils with that margin), which
// pays for the rounding of the fused decode in the traversal. Empty slots: qlo 255, qhi 0 (an inverted box, never entered) and, as a
// belt-and-braces default, the code of slot 0 (entering it would repeat work, never change a result).
struct WideNode {
    float ox, oy, oz, sx;          // origin, scale_x = 2^(e_x + 15)
    int   code[4];                 // same coding as BvhNode children
    uint32_t qlo[3], qhi[3];       // byte c of qlo[k] = child c, axis k
    float sy, sz;
};
#define PRT_EMPTY ((int)0x80000000)

// child / hit code: >= 0 internal node index; < 0 leaf, ~code = (shape_type << 28) | index
#define PRT_LEAF_TYPE(c)  ((uint32_t)(c) >> 28)
#define PRT_LEAF_INDEX(c) ((uint32_t)(c) & 0x0fffffffu)
#define PRT_MAKE_LEAF(type, idx) (((uint32_t)(type) << 28) | (uint32_t)(idx))
#define PRT_NO_HIT 0xffffffffu

struct SceneView {
    const SphereRec* spheres;  int n_spheres;
    const QuadRec*   quads;    int n_quads;
    const float4*    tri_v;    // 3 float4 per leaf slot: {v0, bits(prim_id)} {v1, bits(mat)} {v2, 0}
    const float4*    tri_n;    // 3 float4 per leaf slot: vertex normals
    int              n_tris;
    const MatRec*    mats;     int n_mats;
    const int*       quad_mat; // material per quad (quads' own record is full)
    const BvhNode*   nodes;    int n_prims;
    const WideNode*  wnodes;   // 4-wide quantised nodes: the dense breadth-first copy of the reachable records (k_compact_wide; root = 0) or,
                               // with wide_compact = 0, one record per internal node indexed like `nodes`; NULL = none
    int              root_code; // root: 0 if n_prims>1, leaf code (negative-as-~) if n_prims==1
    float            tri_tmin;
    float sky_r, sky_g, sky_b;
    int              rr_start; // estimator option, default off (-1): Russian roulette from this bounce on
    float            clamp;    // estimator option, default off (+inf): per-path radiance clamp ("firefly clamp")
    uint32_t         one_bits; // 0x3f800000, read from the constant bank so that PRMT keeps its selector as the immediate
    float*           direct;   // != nullptr: a 1-spp single-wave "fresh" render writes each path's contribution straight to accum[3*slot..] (no contribution plane, no resolve)
    int              brute;    // tiny sphere/quad scene: test every primitive in list order, no BVH (see traverse())
    uint32_t         capacity; // slots of the wave's queues / contribution plane (checked builds test every slot against it)
};

// ------------------------------------------------------------------ primitive tests
// Closest-hit candidate bookkeeping: strict < on t, ties -> lowest reference list index
// (primitive.cpp:44 keeps the FIRST primitive on equal distance).
struct Best {
    float t; uint32_t code; int prim_id; float u, v;
};
PRT_DEV void best_init(Best& b) { b.t = __int_as_float(0x7f800000); b.code = PRT_NO_HIT; b.prim_id = 0x7fffffff; b.u = 0.f; b.v = 0.f; }
PRT_DEV void best_update(Best& b, float t, uint32_t code, int prim_id, float u, float v) {
    if (t < b.t || (t == b.t && prim_id < b.prim_id)) { b.t = t; b.code = code; b.prim_id = prim_id; b.u = u; b.v = v; }
}

// Sphere: quadratic exactly as Circle::Intersect (shape.h:157-203) with l = o - c (what
// InvMat*o gives for a translation) and d = normalize(ray.d) (what TransformNormal gives,
// primitive.cpp:30). Returns t (>= tmin) or -1; front = both roots >= tmin.
PRT_DEV float isect_sphere(const SphereRec& s, V3 o, V3 d, bool& front) {
    V3 l = o - xyz(s.a);
    float r = s.a.w, tmin = s.b.x;
    float a = dot(d, d);
    float b = 2.0f * dot(l, d);
    float c = dot(l, l) - r * r;
    float disc = b * b - 4.0f * a * c;
    if (!(disc >= 0.0f)) return -1.0f;
    float sq = sqrtf(disc);
    float t1 = (-b + sq) / (2.0f * a);
    float t2 = (-b - sq) / (2.0f * a);
    if (t1 >= tmin && t2 >= tmin) { front = true; return t1 < t2 ? t1 : t2; }
    front = false;
    if (t1 >= tmin) return t1;
    if (t2 >= tmin) return t2;
    return -1.0f;
}

// Quad: Quad::Intersect (shape.h:213-239) in world space: plane through c with normal n,
// extents along unit u (local x) and v (local z). front = origin on the +n side.
PRT_DEV float isect_quad(const QuadRec& q, V3 o, V3 d, bool& front) {
    V3 n = xyz(q.d);
    V3 oc = o - xyz(q.a);
    float dn = dot(d, n);
    if (fabsf(dn) < 1e-8f) return -1.0f;
    float on = dot(oc, n);
    float t = -on / dn;
    V3 p = oc + d * t;
    float x = dot(p, xyz(q.b));
    float z = dot(p, xyz(q.c));
    float hw = q.a.w, hh = q.b.w, tmin = q.c.w;
    if (t > tmin && (x * x < hw * hw) && (z * z < hh * hh)) { front = on > 0.0f; return t; }
    return -1.0f;
}

// Watertight ray/triangle test (Woop, Benthin, Wald 2013). Per-ray constants:
// kx, ky are the two axes after kz in cyclic order. The paper also swaps them when d[kz] < 0 (to keep the winding). Swapping
// the sheared x and y rows turns every edge function p - q into q - p; NEGATING the x row (-Ax, -Bx, -Cx) does exactly the
// same, bit for bit, signs of zero included ((-Cx)*By - Cy*(-Bx) = q - p), and costs three sign flips with a per-ray mask
// (the sign of sz = 1/d[kz]; sx is stored with the flip applied) instead of a second set of component selections. With the
// axes in cyclic order the selection needs the two predicates of kz only: 2 compares + 18 selects per triangle, where the
// generic comp(v, k) form took 54 instructions (source-level ncu: 7.3 % of k_extend_wide's instructions on the 960 K mesh,
// at 5.6 lanes). U, V, W, det, T equal the swap form's except for the sign of a zero (x - x is +0 either way round), which
// showed as -0 / +0 barycentrics on rays through a vertex: the oracle (prt_oracle.cpp:make_shear) therefore states the same form.
struct RayShear { int kz; float sx, sy, sz; };
PRT_DEV RayShear make_shear(V3 d) {
    RayShear r;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    r.kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    const bool k0 = r.kz == 0, k1 = r.kz == 1;
    float dz = k0 ? d.x : (k1 ? d.y : d.z);
    float dx = k0 ? d.y : (k1 ? d.z : d.x);
    float dy = k0 ? d.z : (k1 ? d.x : d.y);
    r.sz = 1.0f / dz; r.sx = dx * r.sz; r.sy = dy * r.sz;      // one division (as the paper's implementation; every vertex is sheared by the same sx, sy)
    r.sx = __uint_as_float(__float_as_uint(r.sx) ^ (__float_as_uint(r.sz) & 0x80000000u));
    return r;
}
PRT_DEV float comp(V3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }

// Returns t >= tmin or -1; b1,b2 = barycentric weights of v1,v2 (reference naming,
// shape.h:282-284).
PRT_DEV float isect_tri(V3 v0, V3 v1, V3 v2, V3 o, const RayShear& rs, float tmin, float tmax,
                        float& b1, float& b2) {
    V3 A = v0 - o, B = v1 - o, C = v2 - o;
    const bool k0 = rs.kz == 0, k1 = rs.kz == 1;
    float Akz = k0 ? A.x : (k1 ? A.y : A.z), Bkz = k0 ? B.x : (k1 ? B.y : B.z), Ckz = k0 ? C.x : (k1 ? C.y : C.z);
    float Akx = k0 ? A.y : (k1 ? A.z : A.x), Bkx = k0 ? B.y : (k1 ? B.z : B.x), Ckx = k0 ? C.y : (k1 ? C.z : C.x);
    float Aky = k0 ? A.z : (k1 ? A.x : A.y), Bky = k0 ? B.z : (k1 ? B.x : B.y), Cky = k0 ? C.z : (k1 ? C.x : C.y);
    const uint32_t flip = __float_as_uint(rs.sz) & 0x80000000u;            // d[kz] < 0: negate the x row (= the paper's swap)
    Akx = __uint_as_float(__float_as_uint(Akx) ^ flip); Bkx = __uint_as_float(__float_as_uint(Bkx) ^ flip); Ckx = __uint_as_float(__float_as_uint(Ckx) ^ flip);
    float Ax = fmaf(-rs.sx, Akz, Akx), Ay = fmaf(-rs.sy, Akz, Aky);
    float Bx = fmaf(-rs.sx, Bkz, Bkx), By = fmaf(-rs.sy, Bkz, Bky);
    float Cx = fmaf(-rs.sx, Ckz, Ckx), Cy = fmaf(-rs.sy, Ckz, Cky);
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
    // some edge function negative AND some positive = outside (the paper's test; min / max form: 2 three-input min/max + 2 compares)
    if (fminf(fminf(U, V), W) < 0.0f && fmaxf(fmaxf(U, V), W) > 0.0f) return -1.0f;
    float det = (U + V) + W;
    if (det == 0.0f) return -1.0f;
    float Az = rs.sz * Akz, Bz = rs.sz * Bkz, Cz = rs.sz * Ckz;
    float T = fmaf(W, Cz, fmaf(V, Bz, U * Az));
    float rdet = 1.0f / det;                                             // one division per candidate, not three
    float t = T * rdet;
    if (!(t >= tmin) || t > tmax) return -1.0f;
    b1 = V * rdet;
    b2 = W * rdet;
    return t;
}

// ------------------------------------------------------------------ ray/box (robust slab)
struct RayBox { V3 o; V3 inv; V3 pad; };           // pad: 0, except on axes the ray is parallel to (see make_raybox)
PRT_DEV float safe_inv(float d) {
    float a = fabsf(d) < 1e-20f ? copysignf(1e-20f, d) : d;
    return 1.0f / a;
}
// A ray PARALLEL to a slab (direction component 0: inv = +-1e20) that lies exactly in one of the slab's planes gets the
// product 0 there - "leaves at t = 0" - although it stays in the closed slab for every t: an axis-parallel ray through a vertex
// that is the extreme of its box was culled, while the list loop (the reference's algorithm, primitive.cpp:21-59) reports
// the hit (found by tests/test_fuzz_parity.py). On such an axis the two slab products are therefore loosened by P in t
// (isect_box adds it inside the FMA that replaces the multiply: no instruction more), P = twice the largest t any point
// of the scene box can have on the ray's dominant axis. In coordinates that is P * 1e-20: boxes the origin is strictly
// outside of are culled as before. Every other ray has pad = 0 and the products keep their bits.
PRT_DEV RayBox make_raybox(const SceneView& sc, V3 o, V3 d) {
    RayBox r; r.o = o; r.inv = mk(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z)); r.pad = mk(0.f, 0.f, 0.f);
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    if ((ax < 1e-20f || ay < 1e-20f || az < 1e-20f) && sc.n_prims > 1) {         // rare: exactly axis-parallel rays only
        const float4 lmn = __ldg(&sc.nodes[0].lmin), lmx = __ldg(&sc.nodes[0].lmax);
        const float4 rmn = __ldg(&sc.nodes[0].rmin), rmx = __ldg(&sc.nodes[0].rmax);
        const float ex = fmaxf(fabsf(fminf(lmn.x, rmn.x) - o.x), fabsf(fmaxf(lmx.x, rmx.x) - o.x));
        const float ey = fmaxf(fabsf(fminf(lmn.y, rmn.y) - o.y), fabsf(fmaxf(lmx.y, rmx.y) - o.y));
        const float ez = fmaxf(fabsf(fminf(lmn.z, rmn.z) - o.z), fabsf(fmaxf(lmx.z, rmx.z) - o.z));
        float P = (ax >= ay && ax >= az) ? ex * fabsf(r.inv.x) : (ay >= az ? ey * fabsf(r.inv.y) : ez * fabsf(r.inv.z));
        P = fminf(P * 2.0f, 1e30f);
        r.pad = mk(ax < 1e-20f ? copysignf(P, r.inv.x) : 0.f, ay < 1e-20f ? copysignf(P, r.inv.y) : 0.f, az < 1e-20f ? copysignf(P, r.inv.z) : 0.f);
    }
    return r;
}
// Returns entry distance (>=0) if the box [mn,mx] overlaps the ray within [0,tmax], else -1.
// The far side is widened by 1+2^-21 (covers the <=3 ulp error of the three slab products),
// so a primitive that the exact test would report is never culled by rounding.
PRT_DEV float isect_box(const RayBox& r, float4 mn, float4 mx, float tmax) {
    float x0 = fmaf(mn.x - r.o.x, r.inv.x, -r.pad.x), x1 = fmaf(mx.x - r.o.x, r.inv.x, r.pad.x);
    float y0 = fmaf(mn.y - r.o.y, r.inv.y, -r.pad.y), y1 = fmaf(mx.y - r.o.y, r.inv.y, r.pad.y);
    float z0 = fmaf(mn.z - r.o.z, r.inv.z, -r.pad.z), z1 = fmaf(mx.z - r.o.z, r.inv.z, r.pad.z);
    float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax));
    tf = tf * 1.00000047683715820312f;
    return tn <= tf ? tn : -1.0f;
}

// ------------------------------------------------------------------ leaf test dispatch
// d = ray direction, dn = normalize(d) (the sphere test's local direction).
PRT_DEV void test_leaf(const SceneView& sc, uint32_t code, V3 o, V3 d, V3 dn, const RayShear& rs, Best& best) {
    uint32_t type = PRT_LEAF_TYPE(code), idx = PRT_LEAF_INDEX(code);
    if (!PRT_CHECK(type == 2u ? idx < (uint32_t)sc.n_prims : type == 0u ? idx < (uint32_t)sc.n_spheres : (type == 1u && idx < (uint32_t)sc.n_quads), PRT_CHK_NODE)) return;
    if (type == 2u) {
        float4 a = __ldg(&sc.tri_v[3 * idx + 0]);
        float4 b = __ldg(&sc.tri_v[3 * idx + 1]);
        float4 c = __ldg(&sc.tri_v[3 * idx + 2]);
        float b1, b2;
        float t = isect_tri(xyz(a), xyz(b), xyz(c), o, rs, sc.tri_tmin, best.t, b1, b2);
        if (t >= 0.0f) best_update(best, t, code, __float_as_int(a.w), b1, b2);
    } else if (type == 0u) {
        SphereRec s; s.a = __ldg(&sc.spheres[idx].a); s.b = __ldg(&sc.spheres[idx].b);
        bool front;
        float t = isect_sphere(s, o, dn, front);
        if (t >= 0.0f) best_update(best, t, code, __float_as_int(s.b.z), front ? 1.0f : 0.0f, 0.0f);
    } else {
        QuadRec q; q.a = __ldg(&sc.quads[idx].a); q.b = __ldg(&sc.quads[idx].b);
        q.c = __ldg(&sc.quads[idx].c); q.d = __ldg(&sc.quads[idx].d);
        bool front;
        float t = isect_quad(q, o, d, front);
        if (t >= 0.0f) best_update(best, t, code, __float_as_int(q.d.w), front ? 1.0f : 0.0f, 0.0f);
    }
}

// ------------------------------------------------------------------ surface + materials
struct Surface { V3 p; V3 n; bool front; int mat; };

// Rebuild SurfaceInteraction{Position, Normal, IsFrontFace, Material} for a hit record.
template <bool FAST = false>
PRT_DEV Surface make_surface(const SceneView& sc, V3 o, V3 d, float t, uint32_t code, float u, float v) {
    Surface s;
    uint32_t type = PRT_LEAF_TYPE(code), idx = PRT_LEAF_INDEX(code);
    if (type == 2u) {
        float4 a = __ldg(&sc.tri_v[3 * idx + 0]);
        float4 b = __ldg(&sc.tri_v[3 * idx + 1]);
        float4 c = __ldg(&sc.tri_v[3 * idx + 2]);
        float4 na = __ldg(&sc.tri_n[3 * idx + 0]);
        float4 nb = __ldg(&sc.tri_n[3 * idx + 1]);
        float4 nc = __ldg(&sc.tri_n[3 * idx + 2]);
        float b0 = (1.0f - u) - v;                      // (1 - b1 - b2)  shape.h:291
        s.p = (xyz(a) * b0 + xyz(b) * u) + xyz(c) * v;
        V3 n = (xyz(na) * b0 + xyz(nb) * u) + xyz(nc) * v;
        if (dot(n, d) > 0.0f) { n = neg(n); s.front = false; } else s.front = true;   // shape.h:294-302
        s.n = normalize_t<FAST>(n);                     // TransformNormal, primitive.cpp:39
        s.mat = __float_as_int(b.w);
    } else if (type == 0u) {
        float4 a = __ldg(&sc.spheres[idx].a);
        float4 b = __ldg(&sc.spheres[idx].b);
        // local hit point l + dn*t, normal = normalize(local point) (shape.h:194-199), then
        // Position = Mat*p' = p' + c and TransformNormal normalises once more (primitive.cpp:38-39)
        V3 pl = (o - xyz(a)) + normalize_t<FAST>(d) * t;
        V3 n = normalize_t<FAST>(pl);
        s.front = u != 0.0f;
        if (!s.front) n = neg(n);
        s.n = normalize_t<FAST>(n);
        s.p = pl + xyz(a);
        s.mat = __float_as_int(b.y);
    } else {
        float4 qd = __ldg(&sc.quads[idx].d);
        s.p = o + d * t;
        s.front = u != 0.0f;
        s.n = s.front ? xyz(qd) : neg(xyz(qd));
        s.mat = __ldg(&sc.quad_mat[idx]);
    }
    return s;
}

// glm::reflect: I - N * dot(N, I) * 2
PRT_DEV V3 reflect(V3 i, V3 n) { float k = dot(n, i); return i - (n * k) * 2.0f; }

// MaterialHandle::{Emit,Scatter} (material.h:139-161). Returns true when the path
// continues with direction out_d (already passed through Ray::Normalize as
// cpu/renderer.cpp:84 does) and throughput factor atten; emit = Le of the surface.
template <bool FAST = false>
PRT_DEV bool scatter(const MatRec& m, V3 d, const Surface& s, Rng& rng, V3& atten, V3& out_d, V3& emit) {
    int type = __float_as_int(m.b.x);
    emit = mk(0.f, 0.f, 0.f);
    V3 dir;
    if (type == 1) {                                    // Lambertian, material.h:16-31
        V3 ruv = rng_unit_vector<FAST>(rng);
        dir = s.n + ruv;
        if (fabsf(dir.x) < 1e-8f && fabsf(dir.y) < 1e-8f && fabsf(dir.z) < 1e-8f) dir = s.n;
        dir = normalize_t<FAST>(dir);
        atten = mk(m.a.x, m.a.y, m.a.z);
    } else if (type == 2) {                             // Metal, material.h:48-57
        V3 r = normalize_t<FAST>(reflect(d, s.n));
        V3 ruv = rng_unit_vector<FAST>(rng);
        r = r + ruv * m.a.w;
        dir = normalize_t<FAST>(r);
        atten = mk(m.a.x, m.a.y, m.a.z);
        if (!(dot(dir, s.n) > 0.0f)) return false;
    } else if (type == 3) {                             // Dielectric, material.h:76-95
        atten = mk(1.f, 1.f, 1.f);
        float ior = m.a.w;
        float ri = s.front ? div_t<FAST>(1.0f, ior) : ior;
        float cos_t = fminf(dot(neg(d), s.n), 1.0f);
        float sin_t = sqrt_t<FAST>(1.0f - cos_t * cos_t);
        bool cannot = ri * sin_t > 1.0f;
        float r0 = div_t<FAST>(1.0f - ri, 1.0f + ri);
        r0 = r0 * r0;
        float x = 1.0f - cos_t;
        float x5 = ((x * x) * (x * x)) * x;             // pow(1-cos, 5)
        float fres = r0 + (1.0f - r0) * x5;
        // `cannot_refract || fresnel > Random()` short-circuits: no draw when cannot_refract
        bool refl = cannot;
        if (!refl) refl = fres > rng_uniform(rng);
        if (refl) dir = reflect(d, s.n);
        else {                                          // "Reflect" = refract, math.h:45-50
            float c2 = fminf(dot(neg(d), s.n), 1.0f);
            V3 perp = (d + s.n * c2) * ri;
            float k = -sqrt_t<FAST>(fabsf(1.0f - dot(perp, perp)));
            dir = perp + s.n * k;
        }
    } else {                                            // Emissive / invalid: path ends
        if (type == 4) emit = mk(m.a.x, m.a.y, m.a.z);
        return false;
    }
    out_d = normalize_t<FAST>(dir);                             // scatteredRay.Normalize(), cpu/renderer.cpp:84
    return true;
}

// Camera::GetCameraRay (camera.h:104-132)
struct CamView { V3 pos, front, right, up; float tan_half, aspect, w, h; int W, H; FastDiv divW; };
template <bool FAST = false>
PRT_DEV V3 camera_dir(const CamView& c, float px, float py) {
    float ndc_x = div_t<FAST>(px, c.w) * 2.0f - 1.0f;
    float ndc_y = 1.0f - div_t<FAST>(py, c.h) * 2.0f;
    V3 dc = normalize_t<FAST>(mk(ndc_x * c.aspect * c.tan_half, ndc_y * c.tan_half, -1.0f));
    V3 w = (c.right * dc.x + c.up * dc.y) + neg(c.front) * dc.z;
    return normalize_t<FAST>(w);
}

}  // namespace prt
