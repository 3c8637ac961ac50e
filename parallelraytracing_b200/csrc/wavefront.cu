// wavefront.cu — the wavefront path-tracing pipeline over SoA ray queues in HBM:
//   raygen -> [ extend (persistent BVH traversal) -> shade/scatter + compaction ] x depth -> resolve
//
// Behavioural spec = the reference's estimator, identical in its CPU recursion
// (src/backend/cpu/renderer.cpp:59-103), its megakernel loop (cuda_megakernel/renderer.cu:81-119)
// and its wavefront stages (cuda_wavefront/renderer.cu:186-335):
//   L = 0, beta = 1; repeat <= D times: closest hit?  no -> L += beta*sky, stop
//                                       yes -> L += beta*Le; scatter? beta *= atten, continue : stop
// This is a new design, not a port: rays travel IN the queues as 16-byte planes (the reference
// keeps per-pixel state and queues of pixel indices), compaction is by ballot/popc with one
// atomic per warp, extend is a persistent grid sized to the SM count reading the queue length
// from device memory, the RNG is a stateless counter hash (no 48-byte curandState), and the
// per-sample radiance goes to a slot-indexed plane that `resolve` sums in sample order, so the
// accumulated image is deterministic (no float atomics).
#include "prt_internal.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <mutex>

namespace prt {

constexpr int WF_THREADS = 256;
constexpr int STACK_SMEM = 24;    // entries per thread held in shared memory
// Overflow entries in local memory (touched only by the rare deep walk; an untouched local array costs nothing).
// Sized for the worst case, not the typical one: a Karras tree over 30-bit codes + 28-bit index tie-breaks is at most 58
// levels deep; the two-child walk pushes <= 1 entry per level, the 4-wide walk pushes <= 3 per wide node and a wide node
// may descend only ONE binary level (slots [L, RLL, RLR, RR]), so 3 * 58 + 1 sentinel = 175 entries can be needed
// (advisor finding, round 1: 85 entries measured on an adversarial 1 M-triangle scene with the old 64). 256 in total.
constexpr int STACK_TOTAL = 256;
constexpr int STACK_LOCAL = STACK_TOTAL - STACK_SMEM;
#if PRT_CHECKED
static __device__ int g_prt_stack_limit = STACK_TOTAL;   // option check_stack_limit: lets a test prove that the stack check fires
#define PRT_STACK_LIMIT g_prt_stack_limit
#else
#define PRT_STACK_LIMIT STACK_TOTAL
#endif

// queue entry: O = {o.xyz, bits(slot<<5 | bounce)}, D = {d.xyz, bits(path_key)}, B = {beta.rgb, 0}
__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }
// streaming (read-once / write-once) queue traffic: bypass L1, do not pollute it for the BVH
__device__ __forceinline__ float4 ld_stream(const float4* p) { return __ldcs(p); }
__device__ __forceinline__ void st_stream(float4* p, float4 v) { __stcs(p, v); }

// ------------------------------------------------------------------ raygen
template <bool FAST>
__global__ void __launch_bounds__(WF_THREADS)
k_raygen(CamView cam, int pix0, int npix, FastDiv div_npix, int n_paths, uint32_t sample0, uint32_t seed_lo, uint32_t seed_hi,
         float4* __restrict__ qo, float4* __restrict__ qd, float4* __restrict__ qb, uint32_t* __restrict__ qcount,
         const uint32_t* __restrict__ dyn = nullptr) {
    if (dyn) { sample0 = dyn[0]; seed_lo = dyn[1]; seed_hi = dyn[2]; }     // CUDA-graph replay: the per-call values live in memory
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) qcount[0] = (uint32_t)n_paths;
    if (i >= n_paths) return;
    int s_local = (int)fdiv((uint32_t)i, div_npix);
    int pixel = pix0 + (i - s_local * npix);
    int y = (int)fdiv((uint32_t)pixel, cam.divW), x = pixel - y * cam.W;
    V3 d = camera_dir<FAST>(cam, (float)x + 0.5f, (float)y + 0.5f);     // pixel centre, cpu/renderer.cpp:45
    uint32_t key = path_key(seed_lo, seed_hi, (uint32_t)pixel, sample0 + (uint32_t)s_local);
    st_stream(&qo[i], make_float4(cam.pos.x, cam.pos.y, cam.pos.z, __uint_as_float((uint32_t)i << 5)));
    st_stream(&qd[i], make_float4(d.x, d.y, d.z, __uint_as_float(key)));
    st_stream(&qb[i], make_float4(1.f, 1.f, 1.f, 0.f));
}

// Camera rays without a queue: in the staged pipeline the bounce-0 extend and shade kernels compute their ray from its path
// index (same functions, same bits as k_raygen) instead of reading what k_raygen wrote - 48 B written + 80 B read per path
// that never cross HBM (17 of the 50 GB a 64-spp step of the 960 K-triangle scene moved). PRIM: 0 = rays come from the queue,
// 1 = camera rays (bit-exact math), 2 = camera rays (fast_math).
struct PrimaryGen {
    CamView cam; int pix0, npix, n_paths; FastDiv div_npix; uint32_t sample0, seed_lo, seed_hi; const uint32_t* dyn;
};
template <bool FAST>
__device__ __forceinline__ void primary_ray(const PrimaryGen& g, uint32_t i, V3& d, uint32_t& key) {
    uint32_t sample0 = g.sample0, lo = g.seed_lo, hi = g.seed_hi;
    if (g.dyn) { sample0 = g.dyn[0]; lo = g.dyn[1]; hi = g.dyn[2]; }        // CUDA-graph replay: the per-call values live in memory
    const int s_local = (int)fdiv(i, g.div_npix);
    const int pixel = g.pix0 + ((int)i - s_local * g.npix);
    const int y = (int)fdiv((uint32_t)pixel, g.cam.divW), x = pixel - y * g.cam.W;
    d = camera_dir<FAST>(g.cam, (float)x + 0.5f, (float)y + 0.5f);           // pixel centre, cpu/renderer.cpp:45
    key = path_key(lo, hi, (uint32_t)pixel, sample0 + (uint32_t)s_local);
}

// ------------------------------------------------------------------ extend
// Closest hit by stack traversal of the two-child LBVH. One thread per ray; the traversal
// stack lives in shared memory ([entry][thread], conflict-free), overflowing to local memory.
template <bool COUNT, bool BRUTE = false, int SS = STACK_SMEM>
__device__ __forceinline__ Best traverse(const SceneView& sc, V3 o, V3 d, int* __restrict__ sstack,
                                         uint32_t& n_nodes, uint32_t& n_prims) {
    Best best; best_init(best);
    if (sc.n_prims == 0) return best;
    if (BRUTE) {
        // A handful of spheres/quads: the reference's own loop over the list (primitive.cpp:26) is cheaper than
        // walking a BVH and, unlike it, warp-uniform: every lane tests the same record at the same time (the
        // records come in through uniform-address loads). Same closest hit: the (t, prim_id) rule is order-free.
        const V3 dns = sc.n_spheres > 0 ? normalize(d) : d;
        for (int i = 0; i < sc.n_spheres; ++i) {
            SphereRec s; s.a = __ldg(&sc.spheres[i].a); s.b = __ldg(&sc.spheres[i].b);
            bool front;
            float t = isect_sphere(s, o, dns, front);
            if (t >= 0.0f) best_update(best, t, PRT_MAKE_LEAF(0, i), __float_as_int(s.b.z), front ? 1.0f : 0.0f, 0.0f);
        }
        for (int i = 0; i < sc.n_quads; ++i) {
            QuadRec q; q.a = __ldg(&sc.quads[i].a); q.b = __ldg(&sc.quads[i].b);
            q.c = __ldg(&sc.quads[i].c); q.d = __ldg(&sc.quads[i].d);
            bool front;
            float t = isect_quad(q, o, d, front);
            if (t >= 0.0f) best_update(best, t, PRT_MAKE_LEAF(1, i), __float_as_int(q.d.w), front ? 1.0f : 0.0f, 0.0f);
        }
        if (COUNT) n_prims += (uint32_t)sc.n_prims;
        return best;
    }
    RayBox rb = make_raybox(sc, o, d);
    RayShear rs; V3 dn = d;
    if (sc.n_tris > 0) rs = make_shear(d);            // per-ray triangle-test constants: only if there are triangles
    else { rs.kz = 2; rs.sx = rs.sy = rs.sz = 0.f; }
    if (sc.n_spheres > 0) dn = normalize(d);          // Circle's local direction (primitive.cpp:30): only if there are spheres
    int lstack[STACK_TOTAL - SS];
    int sp = 0;
    int cur = sc.root_code;
    for (;;) {
        if (cur >= 0) {
            if (!PRT_CHECK(cur < sc.n_prims - 1, PRT_CHK_NODE)) break;
            const BvhNode* nd = sc.nodes + cur;
            float4 lmin, lmax, rmin, rmax;
            ldg8(&nd->lmin, lmin, lmax); ldg8(&nd->rmin, rmin, rmax);
            if (COUNT) ++n_nodes;
            float tl = isect_box(rb, lmin, lmax, best.t);
            float tr = isect_box(rb, rmin, rmax, best.t);
            int cl = __float_as_int(lmin.w), cr = __float_as_int(lmax.w);
            if (tl >= 0.0f && tr >= 0.0f) {
                int nearc = cl, farc = cr;
                if (tr < tl) { nearc = cr; farc = cl; }
                if (sp < SS) sstack[sp * WF_THREADS] = farc; else if (PRT_CHECK(sp < PRT_STACK_LIMIT, PRT_CHK_STACK)) lstack[sp - SS] = farc; else --sp;   // checked build: drop the entry, keep the walk consistent
                ++sp;
                cur = nearc;
                continue;
            } else if (tl >= 0.0f) { cur = cl; continue; }
            else if (tr >= 0.0f) { cur = cr; continue; }
        } else {
            if (COUNT) ++n_prims;
            test_leaf(sc, (uint32_t)~cur, o, d, dn, rs, best);
        }
        if (sp == 0) break;
        --sp;
        cur = sp < SS ? sstack[sp * WF_THREADS] : lstack[sp - SS];
    }
    return best;
}

template <bool COUNT>
__global__ void __launch_bounds__(WF_THREADS)
k_extend(SceneView sc, const float4* __restrict__ qo, const float4* __restrict__ qd, float4* __restrict__ hits,
         const uint32_t* __restrict__ qcount, int bounce, unsigned long long* __restrict__ stats) {
    __shared__ int sstack[STACK_SMEM * WF_THREADS];
    const uint32_t count = qcount[bounce];
    uint32_t nn = 0, np = 0;
    for (uint32_t i = blockIdx.x * WF_THREADS + threadIdx.x; i < count; i += gridDim.x * WF_THREADS) {
        float4 o4 = ld_stream(&qo[i]), d4 = ld_stream(&qd[i]);
        Best b = traverse<COUNT>(sc, xyz(o4), xyz(d4), sstack + threadIdx.x, nn, np);
        st_stream(&hits[i], make_float4(b.t, __uint_as_float(b.code), b.u, b.v));
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, o); np += __shfl_xor_sync(0xffffffffu, np, o); }
        if ((threadIdx.x & 31) == 0 && (nn | np)) {
            atomicAdd(&stats[PRT_MAX_DEPTH + 0], (unsigned long long)nn);
            atomicAdd(&stats[PRT_MAX_DEPTH + 1], (unsigned long long)np);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&stats[bounce], (unsigned long long)count);
}

// Persistent-thread extend with per-lane dynamic ray fetch (Aila & Laine 2009, "persistent while-while"):
// the grid is sized to what is resident on the SMs; every LANE is a worker that pulls its next ray from
// the queue through a warp-aggregated atomic as soon as its current ray is finished, so a warp never
// idles behind its longest ray (ncu on 1M-10M triangle scenes: 6/32 active lanes without this).
// Inner structure: all lanes first walk internal nodes (postponed leaf), then test the pending leaf;
// when fewer than `refill` lanes are still traversing, the warp returns to the fetch point.
// Order-independent result: (t, prim_id) tie rule makes the closest hit unique, so records are
// bit-identical to k_extend / the CPU oracle; node and leaf counters are identical too.
#define PRT_DONE ((int)0x80000000)
template <bool COUNT, int PRIM = 0>
__global__ void __launch_bounds__(WF_THREADS)
k_extend_pt(SceneView sc, const float4* __restrict__ qo, const float4* __restrict__ qd, float4* __restrict__ hits,
            const uint32_t* __restrict__ qcount, uint32_t* __restrict__ work, int bounce, int refill,
            unsigned long long* __restrict__ stats, PrimaryGen pg) {
    __shared__ int sstack[STACK_SMEM * WF_THREADS];
    const uint32_t count = PRIM ? (uint32_t)pg.n_paths : qcount[bounce];
    const uint32_t lane = threadIdx.x & 31u;
    int* __restrict__ st = sstack + threadIdx.x;
    int lstack[STACK_LOCAL];
    uint32_t nn = 0, np = 0;
    bool have = false, exhausted = (sc.n_prims == 0);
    uint32_t idx = 0; V3 o = mk(0, 0, 0), d = o, dn = o; RayBox rb; RayShear rs; Best best; int cur = PRT_DONE, sp = 0;
    rb.o = o; rb.inv = o; rb.pad = o; rs.kz = 2; rs.sx = rs.sy = rs.sz = 0.f; best_init(best);
    if (sc.n_prims == 0) {          // nothing to hit: every ray is a miss
        for (uint32_t i = blockIdx.x * WF_THREADS + threadIdx.x; i < count; i += gridDim.x * WF_THREADS)
            st_stream(&hits[i], make_float4(best.t, __uint_as_float(best.code), 0.f, 0.f));
    }
    while (!exhausted || __any_sync(0xffffffffu, have)) {
        // ---- fetch: idle lanes take the next rays of the queue (one atomic per warp)
        uint32_t idle = __ballot_sync(0xffffffffu, !have);
        if (idle && !exhausted) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(work, (uint32_t)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!have) {
                idx = base + __popc(idle & ((1u << lane) - 1u));
                if (idx < count) {
                    if (PRIM) { uint32_t key_; o = pg.cam.pos; primary_ray<PRIM == 2>(pg, idx, d, key_); }
                    else { float4 o4 = ld_stream(&qo[idx]), d4 = ld_stream(&qd[idx]); o = xyz(o4); d = xyz(d4); }
                    rb = make_raybox(sc, o, d);
                    if (sc.n_tris > 0) rs = make_shear(d);
                    dn = sc.n_spheres > 0 ? normalize(d) : d;
                    best_init(best);
                    cur = sc.root_code; st[0] = PRT_DONE; sp = 1; have = true;
                }
            }
            exhausted = base + (uint32_t)__popc(idle) >= count;
        }
        // ---- traverse
        if (have) {
            const int thresh = exhausted ? 0 : refill;
            for (;;) {
                while (cur >= 0) {                                     // internal nodes
                    const BvhNode* nd = sc.nodes + cur;
                    float4 lmin, lmax, rmin, rmax;
            ldg8(&nd->lmin, lmin, lmax); ldg8(&nd->rmin, rmin, rmax);
                    if (COUNT) ++nn;
                    float tl = isect_box(rb, lmin, lmax, best.t);
                    float tr = isect_box(rb, rmin, rmax, best.t);
                    int cl = __float_as_int(lmin.w), cr = __float_as_int(lmax.w);
                    // branch-free child selection: near child next, far child (if any) to the stack
                    const bool hl = tl >= 0.0f, hr = tr >= 0.0f;
                    const bool right_first = hr && (!hl || tr < tl);
                    const int nearc = right_first ? cr : cl, farc = right_first ? cl : cr;
                    // Push and pop without divergent branches (ncu: as `if (both) push ... else pop` they were 18 % of the
                    // kernel's warp instructions at 3-4 active lanes). Entry 0 of the stack is a PRT_DONE sentinel, so a
                    // pop never tests for "empty"; the store is predicated and the entry below the top is always read.
                    const bool push = hl && hr, any = hl || hr;
                    if (sp < STACK_SMEM) {
                        int* top = st + sp * WF_THREADS;
                        if (push) *top = farc;
                        const int below = top[-WF_THREADS];
                        cur = any ? nearc : below;
                        sp += push ? 1 : (any ? 0 : -1);
                    } else {                                             // deep stack (rare): spill entries live in local memory
                        if (push && PRT_CHECK(sp < PRT_STACK_LIMIT, PRT_CHK_STACK)) { lstack[sp - STACK_SMEM] = farc; ++sp; }
                        if (any) cur = nearc;
                        else { --sp; cur = sp < STACK_SMEM ? st[sp * WF_THREADS] : lstack[sp - STACK_SMEM]; }
                    }
                    if (__popc(__activemask()) < thresh) break;        // lanes drained away (finished or at a leaf)
                }
                if (cur < 0 && cur != PRT_DONE) {                      // postponed leaf
                    if (COUNT) ++np;
                    test_leaf(sc, (uint32_t)~cur, o, d, dn, rs, best);
                    --sp; cur = sp < STACK_SMEM ? st[sp * WF_THREADS] : lstack[sp - STACK_SMEM];
                }
                if (cur == PRT_DONE) break;
                if (__popc(__activemask()) < thresh) break;            // too few lanes left: go refill
            }
            if (cur == PRT_DONE) {
                st_stream(&hits[idx], make_float4(best.t, __uint_as_float(best.code), best.u, best.v));
                have = false;
            }
        }
    }
    if (COUNT) {
        for (int k = 16; k > 0; k >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, k); np += __shfl_xor_sync(0xffffffffu, np, k); }
        if (lane == 0 && (nn | np)) {
            atomicAdd(&stats[PRT_MAX_DEPTH + 0], (unsigned long long)nn);
            atomicAdd(&stats[PRT_MAX_DEPTH + 1], (unsigned long long)np);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&stats[bounce], (unsigned long long)count);
}

// Persistent-thread extend over the 4-wide quantised nodes (WideNode, prt_device.cuh). Same skeleton as
// k_extend_pt; per node: two 256-bit loads, then for each of the 4 children 6 x (PRMT + FFMA) slab planes.
// Decode: byte q -> g = 1 + q*2^-15 (PRMT into the mantissa of 1.0f), plane t = g*Bp + (A - Bp) with
// Bp = scale*inv = 2^15 * step*inv and A = (origin - o)*inv, i.e. A + q*step*inv in one FFMA. Conservative:
// A is widened by 2^-22 |A| (its two roundings), the quantised box has a 2^-6-step margin (rounding of
// A - Bp is <= 2^-9 step) and the far side gets the same 1+2^-21 factor as the two-child test. The closest hit
// does not depend on any of this (order-free (t, prim_id) rule), only the node count does.
// Resident CTAs per SM the register allocation must allow. Measured (one 16/32-spp wave, Mrays/s): 4 CTAs (63 registers)
// soup 10 M 1156, dragon 960 K 6656, soup 1 M 1643; 5 CTAs (48 registers, 52 B of spills) 1184 / 6593 / 1525 - more warps
// only help the scene that misses L2; 6 CTAs (40) 1150 / 6338; 8 CTAs (32) 595 / 5148; 3 CTAs (80 registers) 1020 / 6213.
#ifndef PRT_EXT_MIN_BLOCKS
#define PRT_EXT_MIN_BLOCKS 4
#endif
template <bool COUNT, int PRIM = 0>
__global__ void __launch_bounds__(WF_THREADS, PRT_EXT_MIN_BLOCKS)
k_extend_wide(SceneView sc, const float4* __restrict__ qo, const float4* __restrict__ qd, float4* __restrict__ hits,
              const uint32_t* __restrict__ qcount, uint32_t* __restrict__ work, int bounce, int refill,
              unsigned long long* __restrict__ stats, PrimaryGen pg) {
    __shared__ int sstack[STACK_SMEM * WF_THREADS];
    const uint32_t count = PRIM ? (uint32_t)pg.n_paths : qcount[bounce];
    const uint32_t lane = threadIdx.x & 31u;
    int* __restrict__ st = sstack + threadIdx.x;
    int lstack[STACK_LOCAL];
    uint32_t nn = 0, np = 0;
    bool have = false, exhausted = false;
    uint32_t idx = 0; V3 o = mk(0, 0, 0), d = o, dn = o, inv = o; RayShear rs; Best best; int cur = PRT_DONE, sp = 0;
    rs.kz = 2; rs.sx = rs.sy = rs.sz = 0.f; best_init(best);
    while (!exhausted || __any_sync(0xffffffffu, have)) {
        uint32_t idle = __ballot_sync(0xffffffffu, !have);
        if (idle && !exhausted) {
            int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(work, (uint32_t)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!have) {
                idx = base + __popc(idle & ((1u << lane) - 1u));
                if (idx < count) {
                    if (PRIM) { uint32_t key_; o = pg.cam.pos; primary_ray<PRIM == 2>(pg, idx, d, key_); }
                    else { float4 o4 = ld_stream(&qo[idx]), d4 = ld_stream(&qd[idx]); o = xyz(o4); d = xyz(d4); }
                    inv = mk(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
                    if (sc.n_tris > 0) rs = make_shear(d);
                    dn = sc.n_spheres > 0 ? normalize(d) : d;
                    best_init(best);
                    cur = sc.root_code; st[0] = PRT_DONE; sp = 1; have = true;
                }
            }
            exhausted = base + (uint32_t)__popc(idle) >= count;
        }
        if (have) {
            const int thresh = exhausted ? 0 : refill;
            const bool nx = inv.x < 0.0f, ny = inv.y < 0.0f, nz = inv.z < 0.0f;
            for (;;) {
                while (cur >= 0) {
                    if (!PRT_CHECK(cur < sc.n_prims - 1, PRT_CHK_NODE)) { cur = PRT_DONE; break; }
                    const WideNode* nd = sc.wnodes + cur;
                    float4 a0, a1, b0, b1;
                    ldg8(nd, a0, a1); ldg8(reinterpret_cast<const char*>(nd) + 32, b0, b1);
                    if (COUNT) ++nn;
                    const float Bx = a0.w * inv.x, By = b1.z * inv.y, Bz = b1.w * inv.z;
                    const float Ax = (a0.x - o.x) * inv.x, Ay = (a0.y - o.y) * inv.y, Az = (a0.z - o.z) * inv.z;
                    const float e = 2.384185791015625e-07f;                                        // 2^-22
                    const float nAx = fmaf(-fabsf(Ax), e, Ax) - Bx, fAx = fmaf(fabsf(Ax), e, Ax) - Bx;
                    const float nAy = fmaf(-fabsf(Ay), e, Ay) - By, fAy = fmaf(fabsf(Ay), e, Ay) - By;
                    const float nAz = fmaf(-fabsf(Az), e, Az) - Bz, fAz = fmaf(fabsf(Az), e, Az) - Bz;
                    const uint32_t lx = __float_as_uint(b0.x), ly = __float_as_uint(b0.y), lz = __float_as_uint(b0.z);
                    const uint32_t hx = __float_as_uint(b0.w), hy = __float_as_uint(b1.x), hz = __float_as_uint(b1.y);
                    const uint32_t wnx = nx ? hx : lx, wfx = nx ? lx : hx;
                    const uint32_t wny = ny ? hy : ly, wfy = ny ? ly : hy;
                    const uint32_t wnz = nz ? hz : lz, wfz = nz ? lz : hz;
                    const int cc[4] = {__float_as_int(a1.x), __float_as_int(a1.y), __float_as_int(a1.z), __float_as_int(a1.w)};
                    uint32_t key[4]; bool hit[4];
                    const uint32_t one = sc.one_bits;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t selb = 0x7604u | ((uint32_t)k << 4);        // bytes: [b.0, a.k, b.2, b.3] of (a = word, b = 1.0f)
                        const float gnx = __uint_as_float(__byte_perm(wnx, one, selb));
                        const float gny = __uint_as_float(__byte_perm(wny, one, selb));
                        const float gnz = __uint_as_float(__byte_perm(wnz, one, selb));
                        const float gfx = __uint_as_float(__byte_perm(wfx, one, selb));
                        const float gfy = __uint_as_float(__byte_perm(wfy, one, selb));
                        const float gfz = __uint_as_float(__byte_perm(wfz, one, selb));
                        float tn = fmaxf(fmaxf(fmaf(gnx, Bx, nAx), fmaf(gny, By, nAy)), fmaxf(fmaf(gnz, Bz, nAz), 0.0f));
                        float tf = fminf(fminf(fmaf(gfx, Bx, fAx), fmaf(gfy, By, fAy)), fminf(fmaf(gfz, Bz, fAz), best.t));
                        tf = tf * 1.00000047683715820312f;
                        hit[k] = tn <= tf;
                        key[k] = hit[k] ? ((__float_as_uint(tn) & ~3u) | (uint32_t)k) : 0xffffffffu;   // tn >= 0: bits order like ints
                    }
                    const uint32_t kmin = min(min(key[0], key[1]), min(key[2], key[3]));
                    const bool any = kmin != 0xffffffffu;
                    const int near = (int)(kmin & 3u);
                    int nearc = cc[3];
                    nearc = key[2] == kmin ? cc[2] : nearc; nearc = key[1] == kmin ? cc[1] : nearc; nearc = key[0] == kmin ? cc[0] : nearc;
                    if (sp <= STACK_SMEM - 3) {                      // common case: predicated pushes, sentinel-terminated pop
                        int t = sp;
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (hit[k] && key[k] > kmin) { st[t * WF_THREADS] = cc[k]; ++t; }   // every hit child but the nearest
                        const int below = st[(t - 1) * WF_THREADS];
                        sp = t - (any ? 0 : 1);
                        cur = any ? nearc : below;
                    } else {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (key[k] != 0xffffffffu && k != near) {
                                if (sp < STACK_SMEM) st[sp * WF_THREADS] = cc[k]; else if (PRT_CHECK(sp < PRT_STACK_LIMIT, PRT_CHK_STACK)) lstack[sp - STACK_SMEM] = cc[k]; else --sp;
                                ++sp;
                            }
                        if (any) cur = nearc;
                        else { --sp; cur = sp < STACK_SMEM ? st[sp * WF_THREADS] : lstack[sp - STACK_SMEM]; }
                    }
                    if (__popc(__activemask()) < thresh) break;
                }
                if (cur < 0 && cur != PRT_DONE) {
                    if (COUNT) ++np;
                    test_leaf(sc, (uint32_t)~cur, o, d, dn, rs, best);
                    --sp; cur = sp < STACK_SMEM ? st[sp * WF_THREADS] : lstack[sp - STACK_SMEM];
                }
                if (cur == PRT_DONE) break;
                if (__popc(__activemask()) < thresh) break;
            }
            if (cur == PRT_DONE) {
                st_stream(&hits[idx], make_float4(best.t, __uint_as_float(best.code), best.u, best.v));
                have = false;
            }
        }
    }
    if (COUNT) {
        for (int k = 16; k > 0; k >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, k); np += __shfl_xor_sync(0xffffffffu, np, k); }
        if (lane == 0 && (nn | np)) {
            atomicAdd(&stats[PRT_MAX_DEPTH + 0], (unsigned long long)nn);
            atomicAdd(&stats[PRT_MAX_DEPTH + 1], (unsigned long long)np);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&stats[bounce], (unsigned long long)count);
}

// The same walk over the 4-wide quantised nodes for ONE ray held by ONE thread (no queue, no refill): k_tail's traversal on
// scenes that have wide nodes. The node step is the one of k_extend_wide above (same decode, same conservative slack,
// nearest hit child first); SS stack entries per thread in shared memory, the rest in local memory.
template <bool COUNT, int SS>
__device__ __forceinline__ Best traverse_wide(const SceneView& sc, V3 o, V3 d, int* __restrict__ st, uint32_t& nn, uint32_t& np) {
    Best best; best_init(best);
    const V3 inv = mk(safe_inv(d.x), safe_inv(d.y), safe_inv(d.z));
    RayShear rs; rs.kz = 2; rs.sx = rs.sy = rs.sz = 0.f;
    if (sc.n_tris > 0) rs = make_shear(d);
    const V3 dn = sc.n_spheres > 0 ? normalize(d) : d;
    const bool nx = inv.x < 0.0f, ny = inv.y < 0.0f, nz = inv.z < 0.0f;
    int lstack[STACK_TOTAL - SS];
    int cur = sc.root_code, sp = 1;
    st[0] = PRT_DONE;
    while (cur != PRT_DONE) {
        if (cur >= 0) {
            if (!PRT_CHECK(cur < sc.n_prims - 1, PRT_CHK_NODE)) break;
            const WideNode* nd = sc.wnodes + cur;
            float4 a0, a1, b0, b1;
            ldg8(nd, a0, a1); ldg8(reinterpret_cast<const char*>(nd) + 32, b0, b1);
            if (COUNT) ++nn;
            const float Bx = a0.w * inv.x, By = b1.z * inv.y, Bz = b1.w * inv.z;
            const float Ax = (a0.x - o.x) * inv.x, Ay = (a0.y - o.y) * inv.y, Az = (a0.z - o.z) * inv.z;
            const float e = 2.384185791015625e-07f;                                        // 2^-22
            const float nAx = fmaf(-fabsf(Ax), e, Ax) - Bx, fAx = fmaf(fabsf(Ax), e, Ax) - Bx;
            const float nAy = fmaf(-fabsf(Ay), e, Ay) - By, fAy = fmaf(fabsf(Ay), e, Ay) - By;
            const float nAz = fmaf(-fabsf(Az), e, Az) - Bz, fAz = fmaf(fabsf(Az), e, Az) - Bz;
            const uint32_t lx = __float_as_uint(b0.x), ly = __float_as_uint(b0.y), lz = __float_as_uint(b0.z);
            const uint32_t hx = __float_as_uint(b0.w), hy = __float_as_uint(b1.x), hz = __float_as_uint(b1.y);
            const uint32_t wnx = nx ? hx : lx, wfx = nx ? lx : hx;
            const uint32_t wny = ny ? hy : ly, wfy = ny ? ly : hy;
            const uint32_t wnz = nz ? hz : lz, wfz = nz ? lz : hz;
            const int cc[4] = {__float_as_int(a1.x), __float_as_int(a1.y), __float_as_int(a1.z), __float_as_int(a1.w)};
            uint32_t key[4];
            const uint32_t one = sc.one_bits;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t selb = 0x7604u | ((uint32_t)k << 4);
                const float gnx = __uint_as_float(__byte_perm(wnx, one, selb)), gny = __uint_as_float(__byte_perm(wny, one, selb));
                const float gnz = __uint_as_float(__byte_perm(wnz, one, selb)), gfx = __uint_as_float(__byte_perm(wfx, one, selb));
                const float gfy = __uint_as_float(__byte_perm(wfy, one, selb)), gfz = __uint_as_float(__byte_perm(wfz, one, selb));
                const float tn = fmaxf(fmaxf(fmaf(gnx, Bx, nAx), fmaf(gny, By, nAy)), fmaxf(fmaf(gnz, Bz, nAz), 0.0f));
                float tf = fminf(fminf(fmaf(gfx, Bx, fAx), fmaf(gfy, By, fAy)), fminf(fmaf(gfz, Bz, fAz), best.t));
                tf = tf * 1.00000047683715820312f;
                key[k] = tn <= tf ? ((__float_as_uint(tn) & ~3u) | (uint32_t)k) : 0xffffffffu;
            }
            const uint32_t kmin = min(min(key[0], key[1]), min(key[2], key[3]));
            int nearc = cc[3];
            nearc = key[2] == kmin ? cc[2] : nearc; nearc = key[1] == kmin ? cc[1] : nearc; nearc = key[0] == kmin ? cc[0] : nearc;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (key[k] != 0xffffffffu && key[k] > kmin) {                         // every hit child but the nearest
                    if (sp < SS) st[sp * WF_THREADS] = cc[k]; else if (PRT_CHECK(sp < PRT_STACK_LIMIT, PRT_CHK_STACK)) lstack[sp - SS] = cc[k]; else --sp;
                    ++sp;
                }
            if (kmin != 0xffffffffu) { cur = nearc; continue; }
        } else {
            if (COUNT) ++np;
            test_leaf(sc, (uint32_t)~cur, o, d, dn, rs, best);
        }
        --sp;
        cur = sp < SS ? st[sp * WF_THREADS] : lstack[sp - SS];
    }
    return best;
}

// Brute force over every primitive = the reference's own algorithm (primitive.cpp:21-59);
// used for parity (`prt_trace_batch(use_bvh=0)`), never on the render path.
__global__ void __launch_bounds__(WF_THREADS)
k_extend_brute(SceneView sc, const uint32_t* __restrict__ sorted_prim, const float4* __restrict__ qo,
               const float4* __restrict__ qd, float4* __restrict__ hits, const uint32_t* __restrict__ qcount, int bounce) {
    const uint32_t count = qcount[bounce];
    for (uint32_t i = blockIdx.x * WF_THREADS + threadIdx.x; i < count; i += gridDim.x * WF_THREADS) {
        float4 o4 = qo[i], d4 = qd[i];
        V3 o = xyz(o4), d = xyz(d4);
        RayShear rs = make_shear(d);
        V3 dn = normalize(d);
        Best b; best_init(b);
        for (int pos = 0; pos < sc.n_prims; ++pos) {
            uint32_t q = sorted_prim[pos];
            uint32_t code = q < (uint32_t)sc.n_spheres ? PRT_MAKE_LEAF(0, q)
                          : q < (uint32_t)(sc.n_spheres + sc.n_quads) ? PRT_MAKE_LEAF(1, q - sc.n_spheres)
                          : PRT_MAKE_LEAF(2, pos);
            test_leaf(sc, code, o, d, dn, rs, b);
        }
        hits[i] = make_float4(b.t, __uint_as_float(b.code), b.u, b.v);
    }
}

// ------------------------------------------------------------------ shade / scatter + compaction
// One path segment after its closest hit is known: L = beta*Le (or beta*sky on a miss), then
// scatter. Returns true when the path continues with (np, ndir, beta*atten).
// Firefly clamp (OFF by default = +inf): a path contributes once, when it ends; that contribution is clamped per channel.
__device__ __forceinline__ V3 clamp_radiance(const SceneView& sc, V3 L) {
    if (!(sc.clamp < __int_as_float(0x7f800000))) return L;             // off: bit-identical to the reference estimator
    return mk(fminf(L.x, sc.clamp), fminf(L.y, sc.clamp), fminf(L.z, sc.clamp));
}

// A path's contribution, written once when the path ends: into the wave's contribution plane (k_resolve sums it into the
// accumulator in sample order), or - prt_render_fresh with one sample and one wave - straight into the accumulator.
__device__ __forceinline__ void put_contrib(const SceneView& sc, float4* __restrict__ contrib, uint32_t slot, V3 L) {
    if (!PRT_CHECK(slot < sc.capacity, PRT_CHK_QUEUE)) return;
    if (sc.direct) { float* a = sc.direct + 3 * (size_t)slot; a[0] = L.x; a[1] = L.y; a[2] = L.z; }
    else st_stream(&contrib[slot], make_float4(L.x, L.y, L.z, 0.f));
}

template <bool FAST = false>
__device__ __forceinline__ bool shade_segment(const SceneView& sc, V3 o, V3 d, uint32_t key, int bounce, int max_depth,
                                              float t, uint32_t code, float hu, float hv, V3& beta, V3& np, V3& ndir, V3& L) {
    if (code == PRT_NO_HIT) {
        L = clamp_radiance(sc, beta * mk(sc.sky_r, sc.sky_g, sc.sky_b));   // cpu/renderer.cpp:99
        return false;
    }
    Surface s = make_surface<FAST>(sc, o, d, t, code, hu, hv);
    MatRec m;
    if (s.mat >= 0 && s.mat < sc.n_mats) { m.a = __ldg(&sc.mats[s.mat].a); m.b = __ldg(&sc.mats[s.mat].b); }
    else { m.a = make_float4(0, 0, 0, 0); m.b = make_float4(0, 0, 0, 0); }
    Rng rng = rng_for_bounce(key, (uint32_t)bounce);
    V3 atten, emit;
    bool sc_ok = scatter<FAST>(m, d, s, rng, atten, ndir, emit);
    L = clamp_radiance(sc, beta * emit);                                // L += emitted (renderer.cpp:74)
    if (sc_ok && bounce + 1 < max_depth) {
        beta = beta * atten;
        // Russian roulette (OFF by default: the reference has none, wavefront.md:98-104 lists it as wanted): survive with
        // p = clamp(max(beta), 0.05, 1), the uniform is the next draw of this bounce's stream, survivors are re-weighted
        if (sc.rr_start >= 0 && bounce >= sc.rr_start) {
            const float p = fmaxf(fminf(fmaxf(fmaxf(beta.x, beta.y), beta.z), 1.0f), 0.05f);
            if (rng_uniform(rng) > p) return false;
            beta = mk(div_t<FAST>(beta.x, p), div_t<FAST>(beta.y, p), div_t<FAST>(beta.z, p));
        }
        np = s.p;
        return true;
    }
    return false;
}

// Queue compaction: ballot/popc, ONE atomic per warp; every lane of the warp must call this.
__device__ __forceinline__ void push_compacted(uint32_t sc_capacity, bool cont, uint32_t lane, uint32_t* __restrict__ next_count,
                                               float4* __restrict__ no, float4* __restrict__ nd, float4* __restrict__ nb,
                                               V3 np, V3 ndir, V3 beta, uint32_t slot_bounce, uint32_t key) {
    uint32_t mask = __ballot_sync(0xffffffffu, cont);
    if (mask) {
        int leader = __ffs(mask) - 1;
        uint32_t b0 = 0;
        if ((int)lane == leader) b0 = atomicAdd(next_count, (uint32_t)__popc(mask));
        b0 = __shfl_sync(0xffffffffu, b0, leader);
        if (cont) {
            uint32_t j = b0 + __popc(mask & ((1u << lane) - 1u));
            if (!PRT_CHECK(j < sc_capacity, PRT_CHK_QUEUE)) return;
            st_stream(&no[j], make_float4(np.x, np.y, np.z, __uint_as_float(slot_bounce)));
            st_stream(&nd[j], make_float4(ndir.x, ndir.y, ndir.z, __uint_as_float(key)));
            st_stream(&nb[j], make_float4(beta.x, beta.y, beta.z, 0.f));
        }
    }
}

// ------------------------------------------------------------------ per-warp shade queue
// Only the rays that HIT something need the expensive part of shading (surface, material, the
// rejection-sampled scatter); on sky-lit scenes that is a minority of a warp's lanes (ncu: the scatter
// code ran at 3.5-7 of 32 lanes). So misses are finished inline (L = beta*sky, one store) and hits are
// appended to a small queue in shared memory owned by the warp; whenever 32 or more are waiting, the warp
// shades 32 of them at full width. No block barrier: the queue is private to the warp. Record = 15 words,
// stored word-major ([word][entry]) so that consecutive lanes touch consecutive banks.
constexpr int SQ_CAP = 64;        // < 32 waiting + <= 32 appended per batch
constexpr int SQ_WORDS = 15;      // o.xyz d.xyz beta.rgb key slot_bounce t code u v
constexpr int SQ_FLOATS = SQ_CAP * SQ_WORDS;
constexpr int WF_WARPS = WF_THREADS / 32;

__device__ __forceinline__ void sq_put(float* __restrict__ q, int e, V3 o, V3 d, V3 beta, uint32_t key, uint32_t slot_bounce,
                                       float t, uint32_t code, float u, float v) {
    q[0 * SQ_CAP + e] = o.x;  q[1 * SQ_CAP + e] = o.y;  q[2 * SQ_CAP + e] = o.z;
    q[3 * SQ_CAP + e] = d.x;  q[4 * SQ_CAP + e] = d.y;  q[5 * SQ_CAP + e] = d.z;
    q[6 * SQ_CAP + e] = beta.x; q[7 * SQ_CAP + e] = beta.y; q[8 * SQ_CAP + e] = beta.z;
    q[9 * SQ_CAP + e] = __uint_as_float(key); q[10 * SQ_CAP + e] = __uint_as_float(slot_bounce);
    q[11 * SQ_CAP + e] = t; q[12 * SQ_CAP + e] = __uint_as_float(code); q[13 * SQ_CAP + e] = u; q[14 * SQ_CAP + e] = v;
}

// Shade queue entries [first, first+n) (n <= 32), one per lane, and push the continuing paths. Whole warp calls.
// A continuing path as it leaves shading; with DEFER the caller pushes it (block-wide, push_block) instead of the warp.
struct ShadeOut { bool cont; V3 np, ndir, beta; uint32_t sb, key; };

template <bool DEFER = false, bool FAST = false>
__device__ __forceinline__ void sq_shade(const SceneView& sc, const float* __restrict__ q, int first, int n, uint32_t lane,
                                         int bounce, int max_depth, float4* __restrict__ contrib, uint32_t* __restrict__ next_count,
                                         float4* __restrict__ no, float4* __restrict__ nd, float4* __restrict__ nb, ShadeOut* out = nullptr) {
    bool cont = false;
    V3 beta = mk(0, 0, 0), np = beta, ndir = beta;
    uint32_t key = 0, sb = 0;
    if ((int)lane < n) {
        const int e = first + (int)lane;
        V3 o = mk(q[0 * SQ_CAP + e], q[1 * SQ_CAP + e], q[2 * SQ_CAP + e]);
        V3 d = mk(q[3 * SQ_CAP + e], q[4 * SQ_CAP + e], q[5 * SQ_CAP + e]);
        beta = mk(q[6 * SQ_CAP + e], q[7 * SQ_CAP + e], q[8 * SQ_CAP + e]);
        key = __float_as_uint(q[9 * SQ_CAP + e]); sb = __float_as_uint(q[10 * SQ_CAP + e]);
        V3 L;
        cont = shade_segment<FAST>(sc, o, d, key, bounce, max_depth, q[11 * SQ_CAP + e], __float_as_uint(q[12 * SQ_CAP + e]),
                             q[13 * SQ_CAP + e], q[14 * SQ_CAP + e], beta, np, ndir, L);
        if (!cont) put_contrib(sc, contrib, sb >> 5, L);
    }
    if (DEFER) { out->cont = cont; out->np = np; out->ndir = ndir; out->beta = beta; out->sb = sb + 1u; out->key = key; }
    else push_compacted(sc.capacity, cont, lane, next_count, no, nd, nb, np, ndir, beta, sb + 1u, key);
}

// Block-wide compaction: ONE atomic on the queue counter per block and round instead of one per warp. All 256 threads
// call it the same number of times. (ncu on the staged k_shade: 57 % of the stall samples sat on the per-warp
// atomicAdd - ~0.6 M same-address atomics per launch at ~2 ns each were the kernel's whole duration.)
__device__ __forceinline__ void push_block(uint32_t sc_capacity, const ShadeOut& o, uint32_t lane, uint32_t* __restrict__ next_count,
                                           float4* __restrict__ no, float4* __restrict__ nd, float4* __restrict__ nb) {
    __shared__ uint32_t wcnt[WF_WARPS];
    __shared__ uint32_t block_base;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t mask = __ballot_sync(0xffffffffu, o.cont);
    if (lane == 0) wcnt[warp] = (uint32_t)__popc(mask);
    __syncthreads();
    uint32_t pre = 0, tot = 0;
#pragma unroll
    for (uint32_t k = 0; k < (uint32_t)WF_WARPS; ++k) { const uint32_t v = wcnt[k]; pre += k < warp ? v : 0u; tot += v; }
    if (threadIdx.x == 0 && tot) block_base = atomicAdd(next_count, tot);
    __syncthreads();
    if (o.cont) {
        const uint32_t j = block_base + pre + (uint32_t)__popc(mask & ((1u << lane) - 1u));
        if (!PRT_CHECK(j < sc_capacity, PRT_CHK_QUEUE)) return;
        st_stream(&no[j], make_float4(o.np.x, o.np.y, o.np.z, __uint_as_float(o.sb)));
        st_stream(&nd[j], make_float4(o.ndir.x, o.ndir.y, o.ndir.z, __uint_as_float(o.key)));
        st_stream(&nb[j], make_float4(o.beta.x, o.beta.y, o.beta.z, 0.f));
    }
}

// One batch of <= 32 segments whose closest hit is known: misses end here, hits go through the queue.
// `live` = this lane holds a segment. Returns the new queue length. Whole warp calls.
template <bool DEFER = false, bool FAST = false>
__device__ __forceinline__ int sq_batch(const SceneView& sc, float* __restrict__ q, int qn, uint32_t lane, bool live,
                                        V3 o, V3 d, V3 beta, uint32_t key, uint32_t slot_bounce, float t, uint32_t code, float u, float v,
                                        int bounce, int max_depth, float4* __restrict__ contrib, uint32_t* __restrict__ next_count,
                                        float4* __restrict__ no, float4* __restrict__ nd, float4* __restrict__ nb, ShadeOut* out = nullptr) {
    const bool hit = live && code != PRT_NO_HIT;
    if (live && !hit) {                                                 // cpu/renderer.cpp:99
        V3 L = clamp_radiance(sc, beta * mk(sc.sky_r, sc.sky_g, sc.sky_b));
        put_contrib(sc, contrib, slot_bounce >> 5, L);
    }
    const uint32_t mask = __ballot_sync(0xffffffffu, hit);
    if (hit) sq_put(q, qn + __popc(mask & ((1u << lane) - 1u)), o, d, beta, key, slot_bounce, t, code, u, v);
    qn += __popc(mask);
    __syncwarp();
    if (qn >= 32) {
        qn -= 32;
        sq_shade<DEFER, FAST>(sc, q, qn, 32, lane, bounce, max_depth, contrib, next_count, no, nd, nb, out);
        __syncwarp();
    }
    return qn;
}


// Resident CTAs per SM asked for k_shade (48 / 56 registers unconstrained = 5 / 4 CTAs). Same box, Mrays/s dragon 960 K 64 spp /
// 809 spheres 64 spp: unconstrained 7 670 / 9 410, 5 CTAs 7 777 / 9 578, 6 CTAs 7 733 / 9 540.
#ifndef PRT_SHADE_MIN_BLOCKS
#define PRT_SHADE_MIN_BLOCKS 5
#endif
template <bool FAST, bool PRIMARY = false>
__global__ void __launch_bounds__(WF_THREADS, PRT_SHADE_MIN_BLOCKS)
k_shade(SceneView sc, const float4* __restrict__ qo, const float4* __restrict__ qd, const float4* __restrict__ qb,
        const float4* __restrict__ hits, float4* __restrict__ no, float4* __restrict__ nd, float4* __restrict__ nb,
        float4* __restrict__ contrib, uint32_t* __restrict__ qcount, int bounce, int max_depth, PrimaryGen pg) {
    __shared__ float sq[WF_WARPS * SQ_FLOATS];
    float* q = sq + (threadIdx.x >> 5) * SQ_FLOATS;
    const uint32_t count = PRIMARY ? (uint32_t)pg.n_paths : qcount[bounce];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * WF_THREADS;
    int qn = 0;
    ShadeOut out;
    for (uint32_t cbase = blockIdx.x * WF_THREADS; cbase < count; cbase += stride) {      // block-uniform trip count
        const uint32_t i = cbase + threadIdx.x;
        float4 o4 = make_float4(0, 0, 0, 0), d4 = o4, b4 = o4, h = o4;
        const bool live = i < count;
        if (live) {
            if (PRIMARY) {                                                  // camera ray recomputed from the path index (see PrimaryGen)
                V3 d; uint32_t key; primary_ray<FAST>(pg, i, d, key);
                o4 = make_float4(pg.cam.pos.x, pg.cam.pos.y, pg.cam.pos.z, __uint_as_float(i << 5));
                d4 = make_float4(d.x, d.y, d.z, __uint_as_float(key)); b4 = make_float4(1.f, 1.f, 1.f, 0.f);
            } else { o4 = ld_stream(&qo[i]); d4 = ld_stream(&qd[i]); b4 = ld_stream(&qb[i]); }
            h = ld_stream(&hits[i]);
        }
        out.cont = false;
        qn = sq_batch<true, FAST>(sc, q, qn, lane, live, xyz(o4), xyz(d4), xyz(b4), __float_as_uint(d4.w), __float_as_uint(o4.w),
                            h.x, __float_as_uint(h.y), h.z, h.w, bounce, max_depth, contrib, &qcount[bounce + 1], no, nd, nb, &out);
        push_block(sc.capacity, out, lane, &qcount[bounce + 1], no, nd, nb);
    }
    out.cont = false;
    if (qn > 0) sq_shade<true, FAST>(sc, q, 0, qn, lane, bounce, max_depth, contrib, &qcount[bounce + 1], no, nd, nb, &out);
    push_block(sc.capacity, out, lane, &qcount[bounce + 1], no, nd, nb);
}

// ------------------------------------------------------------------ fused stages
// The same pipeline with the stage boundaries that only cost HBM traffic removed:
//   k_primary = raygen + extend + shade of bounce 0 (camera rays never touch memory),
//   k_bounce  = extend + shade of bounce b >= 1 (no hit queue); both compact into the next queue.
// Results are bit-identical to the staged kernels (same device functions, same order).
constexpr int FUSED_STACK = 16;   // fused kernels walk trees of <= fused_max_prims primitives; leaves room for the shade queue

// Resident CTAs per SM asked of the register allocator for the two fused kernels (1 = no constraint). The list-loop form of
// k_bounce (BRUTE) compiles to 70 registers unconstrained = 3 CTAs per SM; asked for 4 it takes 64 without a spill. Measured
// on one box, Mrays/s Cornell 256 spp / closed Cornell 64 spp, (bounce, primary): (1,1) 37 527 / 19 696; (3,1) = (4,1)
// 39 205 / 21 200; (5,1) 48 registers + 56 B of spills 38 758 / 21 666; (6,1) 37 978 / 21 347; (4,3) 38 660 / 21 095;
// (4,4) 38 711 / 21 076; (5,3) 38 307 / 21 577; (5,4) 38 456 / 21 551. The BVH-walking form (9 .. 64 primitives, 91
// registers) would spill at 4 and stays unconstrained; k_primary is best left alone.
#ifndef PRT_PRIMARY_MIN_BLOCKS
#define PRT_PRIMARY_MIN_BLOCKS 1
#endif
#ifndef PRT_BOUNCE_MIN_BLOCKS
#define PRT_BOUNCE_MIN_BLOCKS 4
#endif
template <bool COUNT, bool BRUTE, bool FAST>
__global__ void __launch_bounds__(WF_THREADS, PRT_PRIMARY_MIN_BLOCKS)
k_primary(SceneView sc, CamView cam, int pix0, int npix, FastDiv div_npix, int n_paths, uint32_t sample0, uint32_t seed_lo, uint32_t seed_hi,
          float4* __restrict__ no, float4* __restrict__ nd, float4* __restrict__ nb, float4* __restrict__ contrib,
          uint32_t* __restrict__ qcount, int max_depth, unsigned long long* __restrict__ stats, const uint32_t* __restrict__ dyn = nullptr) {
    __shared__ int sstack[BRUTE ? 1 : FUSED_STACK * WF_THREADS];
    __shared__ float sq[WF_WARPS * SQ_FLOATS];
    if (dyn) { sample0 = dyn[0]; seed_lo = dyn[1]; seed_hi = dyn[2]; }     // CUDA-graph replay: the per-call values live in memory
    float* q = sq + (threadIdx.x >> 5) * SQ_FLOATS;
    const uint32_t count = (uint32_t)n_paths;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * WF_THREADS;
    uint32_t nn = 0, npr = 0;
    int qn = 0;
    for (uint32_t base = blockIdx.x * WF_THREADS + (threadIdx.x & ~31u); base < count; base += stride) {
        uint32_t i = base + lane;
        const bool live = i < count;
        V3 d = mk(0, 0, 0);
        uint32_t key = 0;
        Best b; best_init(b);
        if (live) {
            int s_local = (int)fdiv(i, div_npix);
            int pixel = pix0 + ((int)i - s_local * npix);
            int y = (int)fdiv((uint32_t)pixel, cam.divW), x = pixel - y * cam.W;
            d = camera_dir<FAST>(cam, (float)x + 0.5f, (float)y + 0.5f);
            key = path_key(seed_lo, seed_hi, (uint32_t)pixel, sample0 + (uint32_t)s_local);
            b = traverse<COUNT, BRUTE, FUSED_STACK>(sc, cam.pos, d, BRUTE ? sstack : sstack + threadIdx.x, nn, npr);
        }
        qn = sq_batch<false, FAST>(sc, q, qn, lane, live, cam.pos, d, mk(1.f, 1.f, 1.f), key, i << 5, b.t, b.code, b.u, b.v,
                      0, max_depth, contrib, &qcount[1], no, nd, nb);
    }
    if (qn > 0) sq_shade<false, FAST>(sc, q, 0, qn, lane, 0, max_depth, contrib, &qcount[1], no, nd, nb);
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, o); npr += __shfl_xor_sync(0xffffffffu, npr, o); }
        if (lane == 0 && (nn | npr)) {
            atomicAdd(&stats[PRT_MAX_DEPTH + 0], (unsigned long long)nn);
            atomicAdd(&stats[PRT_MAX_DEPTH + 1], (unsigned long long)npr);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { atomicAdd(&stats[0], (unsigned long long)count); qcount[0] = count; }
}

template <bool COUNT, bool BRUTE, bool FAST>
__global__ void __launch_bounds__(WF_THREADS, BRUTE ? PRT_BOUNCE_MIN_BLOCKS : 1)
k_bounce(SceneView sc, const float4* __restrict__ qo, const float4* __restrict__ qd, const float4* __restrict__ qb,
         float4* __restrict__ no, float4* __restrict__ nd, float4* __restrict__ nb, float4* __restrict__ contrib,
         uint32_t* __restrict__ qcount, int bounce, int max_depth, unsigned long long* __restrict__ stats) {
    __shared__ int sstack[BRUTE ? 1 : FUSED_STACK * WF_THREADS];
    __shared__ float sq[WF_WARPS * SQ_FLOATS];
    float* q = sq + (threadIdx.x >> 5) * SQ_FLOATS;
    const uint32_t count = qcount[bounce];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * WF_THREADS;
    uint32_t nn = 0, npr = 0;
    int qn = 0;
    // (fetching the NEXT trip's ray before this trip's work - a third of the stall samples sit on these three loads - cost
    // 12 registers and occupancy: Cornell 38.7 -> 36.9 Grays/s, closed Cornell 21.1 -> 19.3; not in the tree)
    for (uint32_t base = blockIdx.x * WF_THREADS + (threadIdx.x & ~31u); base < count; base += stride) {
        uint32_t i = base + lane;
        const bool live = i < count;
        float4 o4 = make_float4(0, 0, 0, 0), d4 = o4, b4 = o4;
        Best b; best_init(b);
        if (live) {
            o4 = ld_stream(&qo[i]); d4 = ld_stream(&qd[i]); b4 = ld_stream(&qb[i]);
            b = traverse<COUNT, BRUTE, FUSED_STACK>(sc, xyz(o4), xyz(d4), BRUTE ? sstack : sstack + threadIdx.x, nn, npr);
        }
        qn = sq_batch<false, FAST>(sc, q, qn, lane, live, xyz(o4), xyz(d4), xyz(b4), __float_as_uint(d4.w), __float_as_uint(o4.w),
                      b.t, b.code, b.u, b.v, bounce, max_depth, contrib, &qcount[bounce + 1], no, nd, nb);
    }
    if (qn > 0) sq_shade<false, FAST>(sc, q, 0, qn, lane, bounce, max_depth, contrib, &qcount[bounce + 1], no, nd, nb);
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, o); npr += __shfl_xor_sync(0xffffffffu, npr, o); }
        if (lane == 0 && (nn | npr)) {
            atomicAdd(&stats[PRT_MAX_DEPTH + 0], (unsigned long long)nn);
            atomicAdd(&stats[PRT_MAX_DEPTH + 1], (unsigned long long)npr);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&stats[bounce], (unsigned long long)count);
}

// Tail of a SMALL wave. With few rays left, one launch per bounce costs more than the bounce: at 1080p, 1 spp (the
// reference's ProgressiveRender granularity) a k_bounce launch takes ~9 us for < 100 K rays and 3.7 us on an empty
// queue, times 17 bounces at depth 20. k_tail takes the queue of bounce `bounce0` and follows every ray in it to the
// end of its path inside the thread: no queue, no compaction, one launch. Same device functions, same RNG keys, one
// contribution per path => the image and the per-bounce ray counts are bit-identical to the launch-per-bounce form.
template <bool COUNT, bool BRUTE, bool WIDE, bool FAST>
__global__ void __launch_bounds__(WF_THREADS)
k_tail(SceneView sc, const float4* __restrict__ qo, const float4* __restrict__ qd, const float4* __restrict__ qb,
       float4* __restrict__ contrib, const uint32_t* __restrict__ qcount, int bounce0, int max_depth,
       unsigned long long* __restrict__ stats) {
    __shared__ int sstack[BRUTE ? 1 : FUSED_STACK * WF_THREADS];
    __shared__ uint32_t rays_at[PRT_MAX_DEPTH];                        // this block's rays per bounce
    for (int k = threadIdx.x; k < PRT_MAX_DEPTH; k += WF_THREADS) rays_at[k] = 0u;
    __syncthreads();
    const uint32_t count = qcount[bounce0];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * WF_THREADS;
    uint32_t nn = 0, npr = 0;
    for (uint32_t base = blockIdx.x * WF_THREADS + (threadIdx.x & ~31u); base < count; base += stride) {
        const uint32_t i = base + lane;
        bool alive = i < count;
        V3 o = mk(0, 0, 0), d = o, beta = o;
        uint32_t key = 0, slot = 0;
        if (alive) {
            const float4 o4 = ld_stream(&qo[i]), d4 = ld_stream(&qd[i]), b4 = ld_stream(&qb[i]);
            o = xyz(o4); d = xyz(d4); beta = xyz(b4); key = __float_as_uint(d4.w); slot = __float_as_uint(o4.w) >> 5;
        }
        for (int bounce = bounce0; bounce < max_depth; ++bounce) {
            const uint32_t act = __ballot_sync(0xffffffffu, alive);
            if (!act) break;
            if (lane == 0) atomicAdd(&rays_at[bounce], (uint32_t)__popc(act));
            if (alive) {
                const Best b = WIDE ? traverse_wide<COUNT, FUSED_STACK>(sc, o, d, sstack + threadIdx.x, nn, npr)
                                    : traverse<COUNT, BRUTE, FUSED_STACK>(sc, o, d, BRUTE ? sstack : sstack + threadIdx.x, nn, npr);
                V3 np, ndir, L;
                if (shade_segment<FAST>(sc, o, d, key, bounce, max_depth, b.t, b.code, b.u, b.v, beta, np, ndir, L)) { o = np; d = ndir; }
                else { put_contrib(sc, contrib, slot, L); alive = false; }
            }
        }
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, o); npr += __shfl_xor_sync(0xffffffffu, npr, o); }
        if (lane == 0 && (nn | npr)) {
            atomicAdd(&stats[PRT_MAX_DEPTH + 0], (unsigned long long)nn);
            atomicAdd(&stats[PRT_MAX_DEPTH + 1], (unsigned long long)npr);
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < PRT_MAX_DEPTH; k += WF_THREADS)
        if (rays_at[k]) atomicAdd(&stats[k], (unsigned long long)rays_at[k]);
}

// The same tail as a BLOCK-LOCAL WAVEFRONT (fused scenes). k_tail keeps one ray per thread until its path ends: after two
// bounces a third of the lanes are alive (ncu: 9.2 of 32 lanes per instruction, 47 of the 155 us of a one-sample Cornell call).
// Here a block owns a contiguous chunk of the queue and a pool of <= 256 rays in shared memory; every round it (A) finds the
// closest hit of every pooled ray, finishes the misses and packs the hits densely, (B) shades the dense hits at full width,
// (C) packs the surviving paths into the other pool buffer and tops it up from its chunk. No global queue traffic, no launch
// per bounce, warps stay full until the chunk runs dry. Same device functions and RNG keys: image, per-bounce ray counts
// and traversal counters are bit-identical to the launch-per-bounce form.
template <bool COUNT, bool BRUTE, bool FAST>
__global__ void __launch_bounds__(WF_THREADS)
k_tail_pool(SceneView sc, const float4* __restrict__ qo, const float4* __restrict__ qd, const float4* __restrict__ qb,
            float4* __restrict__ contrib, const uint32_t* __restrict__ qcount, int bounce0, int max_depth,
            unsigned long long* __restrict__ stats) {
    __shared__ int sstack[BRUTE ? 1 : FUSED_STACK * WF_THREADS];
    __shared__ float pool[2][11][WF_THREADS];                          // o.xyz d.xyz beta.rgb key slot_bounce, [word][entry]
    __shared__ float hrec[4][WF_THREADS];                              // t, code, u, v of this round's hits, dense
    __shared__ int hsrc[WF_THREADS];                                   // pool entry of each dense hit
    __shared__ uint32_t rays_at[PRT_MAX_DEPTH];
    __shared__ uint32_t s_warp[WF_WARPS];
    const int tid = threadIdx.x;
    const uint32_t lane = tid & 31u, wid = tid >> 5, lt = (1u << lane) - 1u;
    for (int k = tid; k < PRT_MAX_DEPTH; k += WF_THREADS) rays_at[k] = 0u;
    const uint32_t count = qcount[bounce0];
    const uint32_t per = (count + gridDim.x - 1) / gridDim.x;
    uint32_t cursor = min(count, blockIdx.x * per);
    const uint32_t cend = min(count, cursor + per);
    uint32_t nn = 0, npr = 0;
    int cur = 0, pool_n = 0;
    // block-wide exclusive position of a flag + total (two barriers)
    auto pack = [&](bool flag, int& total) -> int {
        const uint32_t m = __ballot_sync(0xffffffffu, flag);
        __syncthreads();                                               // s_warp free again
        if (lane == 0) s_warp[wid] = (uint32_t)__popc(m);
        __syncthreads();
        int base = 0; total = 0;
#pragma unroll
        for (int k = 0; k < WF_WARPS; ++k) { const int v = (int)s_warp[k]; base += k < (int)wid ? v : 0; total += v; }
        return base + __popc(m & lt);
    };
    for (;;) {
        // ---- (C') top the pool up from this block's chunk of the queue
        const int take = min((int)(cend - cursor), WF_THREADS - pool_n);
        if (tid < take) {
            const float4 o4 = ld_stream(&qo[cursor + tid]), d4 = ld_stream(&qd[cursor + tid]), b4 = ld_stream(&qb[cursor + tid]);
            const int e = pool_n + tid;
            pool[cur][0][e] = o4.x; pool[cur][1][e] = o4.y; pool[cur][2][e] = o4.z;
            pool[cur][3][e] = d4.x; pool[cur][4][e] = d4.y; pool[cur][5][e] = d4.z;
            pool[cur][6][e] = b4.x; pool[cur][7][e] = b4.y; pool[cur][8][e] = b4.z;
            pool[cur][9][e] = d4.w; pool[cur][10][e] = o4.w;
        }
        cursor += (uint32_t)take; pool_n += take;
        __syncthreads();
        if (pool_n == 0) break;
        if (cursor == cend && pool_n <= 32) {
            // the chunk is used up and what is left fits one warp: the remaining bounces are a latency chain, and a round
            // of the pool costs five block barriers - warp 0 follows those few paths to their end in-thread (as k_tail does)
            if (wid == 0 && (int)lane < pool_n) {
                V3 o = mk(pool[cur][0][lane], pool[cur][1][lane], pool[cur][2][lane]);
                V3 d = mk(pool[cur][3][lane], pool[cur][4][lane], pool[cur][5][lane]);
                V3 beta = mk(pool[cur][6][lane], pool[cur][7][lane], pool[cur][8][lane]);
                const uint32_t key = __float_as_uint(pool[cur][9][lane]), sb0 = __float_as_uint(pool[cur][10][lane]);
                for (int bounce = (int)(sb0 & 31u); bounce < max_depth; ++bounce) {
                    atomicAdd(&rays_at[bounce], 1u);
                    const Best b = traverse<COUNT, BRUTE, FUSED_STACK>(sc, o, d, BRUTE ? sstack : sstack + tid, nn, npr);
                    V3 np, ndir, L;
                    if (shade_segment<FAST>(sc, o, d, key, bounce, max_depth, b.t, b.code, b.u, b.v, beta, np, ndir, L)) { o = np; d = ndir; }
                    else { put_contrib(sc, contrib, sb0 >> 5, L); break; }
                }
            }
            break;
        }
        // ---- (A) closest hit of every pooled ray; misses end here
        bool hit = false;
        Best b; best_init(b);
        if (tid < pool_n) {
            const V3 o = mk(pool[cur][0][tid], pool[cur][1][tid], pool[cur][2][tid]);
            const V3 d = mk(pool[cur][3][tid], pool[cur][4][tid], pool[cur][5][tid]);
            const uint32_t sb = __float_as_uint(pool[cur][10][tid]);
            atomicAdd(&rays_at[sb & 31u], 1u);
            b = traverse<COUNT, BRUTE, FUSED_STACK>(sc, o, d, BRUTE ? sstack : sstack + tid, nn, npr);
            hit = b.code != PRT_NO_HIT;
            if (!hit) {
                const V3 beta = mk(pool[cur][6][tid], pool[cur][7][tid], pool[cur][8][tid]);
                put_contrib(sc, contrib, sb >> 5, clamp_radiance(sc, beta * mk(sc.sky_r, sc.sky_g, sc.sky_b)));   // cpu/renderer.cpp:99
            }
        }
        int n_hits;
        const int hp = pack(hit, n_hits);
        if (hit) { hrec[0][hp] = b.t; hrec[1][hp] = __uint_as_float(b.code); hrec[2][hp] = b.u; hrec[3][hp] = b.v; hsrc[hp] = tid; }
        __syncthreads();
        // ---- (B) shade the dense hits
        bool cont = false;
        V3 np = mk(0, 0, 0), ndir = np, beta = np; uint32_t key = 0, sb = 0;
        if (tid < n_hits) {
            const int e = hsrc[tid];
            const V3 o = mk(pool[cur][0][e], pool[cur][1][e], pool[cur][2][e]);
            const V3 d = mk(pool[cur][3][e], pool[cur][4][e], pool[cur][5][e]);
            beta = mk(pool[cur][6][e], pool[cur][7][e], pool[cur][8][e]);
            key = __float_as_uint(pool[cur][9][e]); sb = __float_as_uint(pool[cur][10][e]);
            V3 L;
            cont = shade_segment<FAST>(sc, o, d, key, (int)(sb & 31u), max_depth, hrec[0][tid], __float_as_uint(hrec[1][tid]), hrec[2][tid],
                                       hrec[3][tid], beta, np, ndir, L);
            if (!cont) put_contrib(sc, contrib, sb >> 5, L);
        }
        // ---- (C) survivors into the other pool buffer
        int n_cont;
        const int cp = pack(cont, n_cont);
        if (cont) {
            const int nx = cur ^ 1;
            pool[nx][0][cp] = np.x; pool[nx][1][cp] = np.y; pool[nx][2][cp] = np.z;
            pool[nx][3][cp] = ndir.x; pool[nx][4][cp] = ndir.y; pool[nx][5][cp] = ndir.z;
            pool[nx][6][cp] = beta.x; pool[nx][7][cp] = beta.y; pool[nx][8][cp] = beta.z;
            pool[nx][9][cp] = __uint_as_float(key); pool[nx][10][cp] = __uint_as_float(sb + 1u);
        }
        cur ^= 1; pool_n = n_cont;
        __syncthreads();
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, o); npr += __shfl_xor_sync(0xffffffffu, npr, o); }
        if (lane == 0 && (nn | npr)) {
            atomicAdd(&stats[PRT_MAX_DEPTH + 0], (unsigned long long)nn);
            atomicAdd(&stats[PRT_MAX_DEPTH + 1], (unsigned long long)npr);
        }
    }
    __syncthreads();
    for (int k = tid; k < PRT_MAX_DEPTH; k += WF_THREADS)
        if (rays_at[k]) atomicAdd(&stats[k], (unsigned long long)rays_at[k]);
}

// ------------------------------------------------------------------ resolve (accumulate)
// accum[3p+c] += sum over the wave's samples, in sample order (Film::AddSample, film.cu:37-55)
__global__ void __launch_bounds__(WF_THREADS)
k_resolve(const float4* __restrict__ contrib, int pix0, int npix, int n_samples, float* __restrict__ accum) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    float* a = accum + 3 * (size_t)(pix0 + i);
    float r = a[0], g = a[1], b = a[2];
    for (int s = 0; s < n_samples; ++s) {
        float4 c = ld_stream(&contrib[(size_t)s * npix + i]);
        r += c.x; g += c.y; b += c.z;
    }
    a[0] = r; a[1] = g; a[2] = b;
}

// ------------------------------------------------------------------ film display resolve
// Film::UpdateDisplay (film.cu:134-194): avg -> Reinhard -> pow(1/gamma) -> u8, alpha 255.
__global__ void k_display(const float* __restrict__ accum, int npix, float inv_w, float has_w, float exposure,
                          float inv_gamma, uint8_t* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    uint8_t px[4];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float v = 0.0f;
        if (has_w > 0.0f) {
            v = accum[3 * (size_t)i + c] * inv_w;
            float x = v * exposure;
            x = x / (1.0f + x);
            v = (float)pow((double)x, (double)inv_gamma);
        }
        v = fmaxf(0.0f, fminf(1.0f, v));
        px[c] = (uint8_t)(v * 255.0f + 0.5f);
    }
    px[3] = 255;
    *reinterpret_cast<uchar4*>(out + 4 * (size_t)i) = make_uchar4(px[0], px[1], px[2], px[3]);
}

// ------------------------------------------------------------------ hit records for the parity batch
__global__ void k_hit_records(SceneView sc, const float4* __restrict__ qo, const float4* __restrict__ qd,
                              const float4* __restrict__ hits, int n, prt_hit* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 h = hits[i];
    uint32_t code = __float_as_uint(h.y);
    prt_hit r;
    if (code == PRT_NO_HIT) {
        r.t = h.x; r.prim_id = -1; r.b1 = r.b2 = 0.f; r.px = r.py = r.pz = 0.f; r.nx = r.ny = r.nz = 0.f;
        r.front = 0; r.material = -1;
    } else {
        V3 o = xyz(qo[i]), d = xyz(qd[i]);
        Surface s = make_surface(sc, o, d, h.x, code, h.z, h.w);
        uint32_t type = PRT_LEAF_TYPE(code), idx = PRT_LEAF_INDEX(code);
        int pid = type == 2u ? __float_as_int(sc.tri_v[3 * idx].w)
                : type == 0u ? __float_as_int(sc.spheres[idx].b.z) : __float_as_int(sc.quads[idx].d.w);
        r.t = h.x; r.prim_id = pid;
        r.b1 = type == 2u ? h.z : 0.f; r.b2 = type == 2u ? h.w : 0.f;
        r.px = s.p.x; r.py = s.p.y; r.pz = s.p.z; r.nx = s.n.x; r.ny = s.n.y; r.nz = s.n.z;
        r.front = s.front ? 1 : 0; r.material = s.mat;
    }
    out[i] = r;
}

__global__ void k_load_rays(const prt_ray* __restrict__ rays, int n, float4* __restrict__ qo, float4* __restrict__ qd,
                            uint32_t* __restrict__ qcount) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) qcount[0] = (uint32_t)n;
    if (i >= n) return;
    prt_ray r = rays[i];
    qo[i] = make_float4(r.ox, r.oy, r.oz, 0.f);
    qd[i] = make_float4(r.dx, r.dy, r.dz, 0.f);
}

// ------------------------------------------------------------------ host side
SceneView scene_view(const Ctx* c) {
    SceneView v{};
    v.spheres = (const SphereRec*)c->spheres.p; v.n_spheres = c->n_spheres;
    v.quads = (const QuadRec*)c->quads.p; v.n_quads = c->n_quads;
    v.tri_v = (const float4*)c->tri_v.p; v.tri_n = (const float4*)c->tri_n.p; v.n_tris = c->n_tris;
    v.mats = (const MatRec*)c->mats.p; v.n_mats = c->n_mats;
    v.quad_mat = (const int*)c->quad_mat.p;
    v.nodes = (const BvhNode*)c->nodes.p; v.n_prims = c->n_prims;
    v.wnodes = c->wide_built ? (const WideNode*)(c->wide_compacted ? c->wnodes_c.p : c->wnodes.p) : nullptr;
    v.root_code = c->root_code;
    v.tri_tmin = 0.001f;                                  // kShapeRayTMin, shape.h:128
    v.sky_r = c->sky[0]; v.sky_g = c->sky[1]; v.sky_b = c->sky[2];
    v.one_bits = 0x3f800000u;
    v.rr_start = c->rr_start; v.clamp = c->firefly_clamp > 0.f ? c->firefly_clamp : INFINITY;
    v.brute = 0;                                          // only wavefront_render's fused path turns it on
    v.capacity = 0xffffffffu;
    v.direct = nullptr;
    return v;
}

static CamView cam_view(const prt_camera& k) {
    CamView v;
    v.pos = {k.pos[0], k.pos[1], k.pos[2]};
    v.front = {k.front[0], k.front[1], k.front[2]};
    v.right = {k.right[0], k.right[1], k.right[2]};
    v.up = {k.up[0], k.up[1], k.up[2]};
    v.tan_half = k.tan_half_fov;
    v.w = (float)k.width; v.h = (float)k.height;
    v.aspect = v.w / v.h;                                 // m_Width / m_Height, camera.h:112
    v.W = k.width; v.H = k.height;
    v.divW = make_fastdiv((uint32_t)k.width);
    return v;
}

static int ensure_wave_buffers(Ctx* c, int64_t paths, bool second_set = false) {
    int rc;
    size_t b16 = (size_t)paths * 16;
    for (int k = 0; k < 2; ++k)
        if ((rc = dev_reserve(c, c->q_o[k], b16)) || (rc = dev_reserve(c, c->q_d[k], b16)) ||
            (rc = dev_reserve(c, c->q_b[k], b16))) return rc;
    if ((rc = dev_reserve(c, c->hits, b16)) || (rc = dev_reserve(c, c->contrib, b16))) return rc;
    if ((rc = dev_reserve(c, c->qcount, 2 * (PRT_MAX_DEPTH + 2) * 4))) return rc;   // queue lengths + per-launch work counters
    if ((rc = dev_reserve(c, c->stats_dev, (PRT_MAX_DEPTH + 4) * 8))) return rc;
    if (second_set) {
        for (int k = 0; k < 2; ++k)
            if ((rc = dev_reserve(c, c->q2_o[k], b16)) || (rc = dev_reserve(c, c->q2_d[k], b16)) ||
                (rc = dev_reserve(c, c->q2_b[k], b16))) return rc;
        if ((rc = dev_reserve(c, c->hits2, b16)) || (rc = dev_reserve(c, c->contrib2, b16))) return rc;
        if ((rc = dev_reserve(c, c->qcount2, 2 * (PRT_MAX_DEPTH + 2) * 4))) return rc;
        if (!c->stream2) {
            PRT_CUDA(c, cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking));
            for (cudaEvent_t* e : {&c->ev_res[0], &c->ev_res[1], &c->ev_join})
                PRT_CUDA(c, cudaEventCreateWithFlags(e, cudaEventDisableTiming));
        }
    }
    return PRT_OK;
}

int wavefront_render(Ctx* c, int spp, int max_depth, uint64_t seed, int sample_begin, int64_t pixel_begin, int64_t pixel_count, bool fresh, bool defer_finish) {
    const int64_t P = pixel_count;                                   // pixels [pixel_begin, pixel_begin + P) of the frame
    int64_t cap = c->wave_paths_opt > 0 ? c->wave_paths_opt : (int64_t)64 << 20;   // 64 Mi paths: 8.6 GB of queues per set
    cap = std::min<int64_t>(cap, (int64_t)1 << 26);                  // slot must fit 27 bits
    // wave = pixel range [p0, p0+npix) x S samples
    int64_t npix_wave = std::min<int64_t>(P, cap);
    int S_wave = (int)std::max<int64_t>(1, std::min<int64_t>(spp, cap / npix_wave));
    const int64_t n_waves = ((spp + S_wave - 1) / S_wave) * ((P + npix_wave - 1) / npix_wave);
    const int nsets = (c->overlap && n_waves > 1) ? 2 : 1;           // two waves in flight on two streams
    int rc;
    if ((rc = ensure_wave_buffers(c, npix_wave * S_wave, nsets > 1))) return rc;
    c->wave_paths = npix_wave * S_wave;
    cudaStream_t sts[2] = {c->stream, nsets > 1 ? c->stream2 : c->stream};

    SceneView sc = scene_view(c);
    sc.capacity = (uint32_t)(npix_wave * S_wave);                     // slots of one wave's queues and contribution plane
    CamView cam = cam_view(c->cam);
    // fresh = "clear, then render": with ONE sample in ONE wave every pixel of the range gets exactly one contribution, so the
    // paths write it straight into the accumulator (no memset, no contribution plane, no resolve); otherwise clear first
    const bool direct = fresh && n_waves == 1 && spp == 1;
    if (fresh && !direct) PRT_CUDA(c, cudaMemsetAsync(c->accum + 3 * pixel_begin, 0, (size_t)P * 12, c->stream));
    if (direct) sc.direct = c->accum + 3 * pixel_begin;
    unsigned long long* dstats = (unsigned long long*)c->stats_dev.p;
    const bool brute = c->n_tris == 0 && c->n_prims <= c->brute_max_prims;   // fused kernels only
    sc.brute = brute ? 1 : 0;
    const int grid_p = c->num_sms * c->ctas_per_sm;
    const uint32_t seed_lo = (uint32_t)seed, seed_hi = (uint32_t)(seed >> 32);
    const bool fused = c->fused < 0 ? (c->n_prims <= c->fused_max_prims) : (c->fused != 0);
    const bool timing_opt = c->stage_timing != 0;
    const bool fastm = c->fast_math != 0;                              // tolerance-mode float math on the shading side (prt_device.cuh)
    const bool wide = sc.wnodes != nullptr && c->n_prims > 1;          // staged extend walks the 4-wide quantised nodes
    auto k_ext = wide ? (c->counters_on ? k_extend_wide<true, 0> : k_extend_wide<false, 0>)
                      : (c->counters_on ? k_extend_pt<true, 0> : k_extend_pt<false, 0>);
    // bounce 0 of the staged pipeline: camera rays are computed in the kernels (no k_raygen, no queue traffic for them)
    auto k_ext0 = fastm ? (wide ? (c->counters_on ? k_extend_wide<true, 2> : k_extend_wide<false, 2>)
                                : (c->counters_on ? k_extend_pt<true, 2> : k_extend_pt<false, 2>))
                        : (wide ? (c->counters_on ? k_extend_wide<true, 1> : k_extend_wide<false, 1>)
                                : (c->counters_on ? k_extend_pt<true, 1> : k_extend_pt<false, 1>));
    int grid_pt = grid_p;
    {   // persistent extend: exactly the CTAs that are resident at once
        int per_sm = 0;
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ext, WF_THREADS, 0);
        if (e == cudaSuccess && per_sm > 0) grid_pt = c->num_sms * per_sm;
    }

    // per-stage device timing: pooled events, one pair per kernel of the two hot stages
    size_t ev_used = 0;
    std::vector<std::pair<size_t, int>>& spans = c->pend_spans;      // (first event index, stage)
    spans.clear();
    auto stamp = [&](cudaStream_t s) -> size_t {
        if (ev_used == c->ev_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); c->ev_pool.push_back(e); }
        cudaEventRecord(c->ev_pool[ev_used], s);
        return ev_used++;
    };

    // ---- the reference's call pattern: ONE sample per call (main.cpp:513). Such a call is 3-20 short launches + two memsets +
    // the counter read-back; issued one by one they cost 26 us of host time against 155 us of device time (Cornell 1080p).
    // The whole enqueue is captured ONCE into a CUDA graph per (scene, camera, depth, tail start) and replayed with one
    // cudaGraphLaunch; what changes per call (first sample index, seed) reaches k_primary / k_raygen through three words that
    // a memcpy node of the graph copies from pinned host memory at replay time.
    if (!c->stats_host) PRT_CUDA(c, cudaHostAlloc((void**)&c->stats_host, (PRT_MAX_DEPTH + 4) * 8, cudaHostAllocDefault));
    const bool use_graph = direct && c->use_graph && !c->sort_rays;
    const uint32_t* dyn = nullptr;
    bool capturing = false;
    GraphKey gkey{};
    if (use_graph) {
        if (!c->dyn_host) PRT_CUDA(c, cudaHostAlloc((void**)&c->dyn_host, 16, cudaHostAllocDefault));
        if ((rc = dev_reserve(c, c->dyn_dev, 16))) return rc;
        c->dyn_host[0] = (uint32_t)sample_begin; c->dyn_host[1] = seed_lo; c->dyn_host[2] = seed_hi;
        int tb = (c->tail_bounce > 0 && P <= c->tail_max_paths) ? c->tail_bounce : max_depth;        // same rule as in the wave loop
        if (tb < max_depth && c->tail_adaptive) {
            const unsigned long long h0 = c->hint_rays[0];
            tb = max_depth;
            for (int b = 1; b < max_depth; ++b) {
                const double expect = h0 ? (double)P * (double)c->hint_rays[b] / (double)h0 : (double)P;
                if (expect <= (double)(fused ? c->tail_rays : c->tail_rays_staged)) { tb = b; break; }
            }
        }
        gkey = GraphKey{c->graph_gen, pixel_begin, P, max_depth, tb, c->counters_on, (void*)c->accum};
        for (GraphEntry& g : c->graphs)
            if (g.key == gkey) {                                      // replay
                PRT_CUDA(c, cudaEventRecord(c->ev0, c->stream));
                PRT_CUDA(c, cudaGraphLaunch(g.exec, c->stream));
                PRT_CUDA(c, cudaEventRecord(c->ev1, c->stream));
                spans.clear();
                c->pend_launches = g.launches; c->pend_paths = (uint64_t)P; c->pend_active = true;
                c->accum_spp = spp;
                if (defer_finish) return PRT_OK;
                return wavefront_finish(c);
            }
        PRT_CUDA(c, cudaEventRecord(c->ev0, c->stream));
        PRT_CUDA(c, cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
        capturing = true;
        dyn = (const uint32_t*)c->dyn_dev.p;
    }
    struct CaptureGuard {                                             // an error return below must not leave the stream capturing
        Ctx* c; bool* on;
        ~CaptureGuard() { if (*on) { cudaGraph_t g = nullptr; cudaStreamEndCapture(c->stream, &g); if (g) cudaGraphDestroy(g); } }
    } guard{c, &capturing};
    if (capturing) PRT_CUDA(c, cudaMemcpyAsync(c->dyn_dev.p, c->dyn_host, 12, cudaMemcpyHostToDevice, c->stream));
    const bool timing = timing_opt && !capturing && !(direct && n_waves == 1);   // event pairs per kernel: measurable at one sample per call

    PRT_CUDA(c, cudaMemsetAsync(dstats, 0, (PRT_MAX_DEPTH + 4) * 8, c->stream));
    if (!capturing) PRT_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    if (nsets > 1) PRT_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev0, 0));
    int launches = 0;
    uint64_t paths = 0;
    int64_t wave = 0;
    for (int s0 = 0; s0 < spp; s0 += S_wave) {
        const int S = std::min(S_wave, spp - s0);
        for (int64_t p0 = pixel_begin; p0 < pixel_begin + P; p0 += npix_wave, ++wave) {
            const int set = (int)(wave % nsets);
            cudaStream_t st = sts[set];
            const int npix = (int)std::min<int64_t>(npix_wave, pixel_begin + P - p0);
            const int n_paths = npix * S;
            const FastDiv div_npix = make_fastdiv((uint32_t)npix);
            paths += (uint64_t)n_paths;
            uint32_t* qcount = (uint32_t*)(set ? c->qcount2.p : c->qcount.p);
            PRT_CUDA(c, cudaMemsetAsync(qcount, 0, 2 * (PRT_MAX_DEPTH + 2) * 4, st));
            const int grid = (int)std::min<int64_t>(grid_p, (n_paths + WF_THREADS - 1) / WF_THREADS);
            DevBuf *bo = set ? c->q2_o : c->q_o, *bd = set ? c->q2_d : c->q_d, *bb = set ? c->q2_b : c->q_b;
            float4 *qo[2] = {(float4*)bo[0].p, (float4*)bo[1].p}, *qd[2] = {(float4*)bd[0].p, (float4*)bd[1].p},
                   *qb[2] = {(float4*)bb[0].p, (float4*)bb[1].p};
            float4 *hits = (float4*)(set ? c->hits2.p : c->hits.p), *contrib = (float4*)(set ? c->contrib2.p : c->contrib.p);
            int cur = 0;
            // small waves: launch-per-bounce up to tail_bounce, then ONE kernel finishes the paths in-thread
            int tail_b = (c->tail_bounce > 0 && n_paths <= c->tail_max_paths) ? c->tail_bounce : max_depth;
            if (tail_b < max_depth && c->tail_adaptive) {
                const unsigned long long p0 = c->hint_rays[0];
                tail_b = max_depth;
                for (int b = 1; b < max_depth; ++b) {          // expected queue length at bounce b
                    const double expect = p0 ? (double)n_paths * (double)c->hint_rays[b] / (double)p0 : (double)n_paths;
                    if (expect <= (double)(fused ? c->tail_rays : c->tail_rays_staged)) { tail_b = b; break; }
                }
            }
            if (fused) {
                if (timing) spans.push_back({stamp(st), 2});
                {
                    auto kp = fastm ? (c->counters_on ? (brute ? k_primary<true, true, true> : k_primary<true, false, true>)
                                                      : (brute ? k_primary<false, true, true> : k_primary<false, false, true>))
                                    : (c->counters_on ? (brute ? k_primary<true, true, false> : k_primary<true, false, false>)
                                                      : (brute ? k_primary<false, true, false> : k_primary<false, false, false>));
                    kp<<<grid, WF_THREADS, 0, st>>>(sc, cam, (int)p0, npix, div_npix, n_paths, (uint32_t)(sample_begin + s0), seed_lo,
                                                    seed_hi, qo[1], qd[1], qb[1], contrib, qcount, max_depth, dstats, dyn);
                }
                if (timing) stamp(st);
                ++launches;
                cur = 1;
                for (int b = 1; b < std::min(max_depth, tail_b); ++b) {
                    if (timing) spans.push_back({stamp(st), 0});
                    {
                        auto kb = fastm ? (c->counters_on ? (brute ? k_bounce<true, true, true> : k_bounce<true, false, true>)
                                                          : (brute ? k_bounce<false, true, true> : k_bounce<false, false, true>))
                                        : (c->counters_on ? (brute ? k_bounce<true, true, false> : k_bounce<true, false, false>)
                                                          : (brute ? k_bounce<false, true, false> : k_bounce<false, false, false>));
                        kb<<<grid, WF_THREADS, 0, st>>>(sc, qo[cur], qd[cur], qb[cur], qo[cur ^ 1], qd[cur ^ 1], qb[cur ^ 1],
                                                        contrib, qcount, b, max_depth, dstats);
                    }
                    if (timing) stamp(st);
                    ++launches;
                    cur ^= 1;
                }
            } else {
                const PrimaryGen pg{cam, (int)p0, npix, n_paths, div_npix, (uint32_t)(sample_begin + s0), seed_lo, seed_hi, dyn};
                const bool gen0 = c->persistent != 0;                  // the persistent extend kernels generate camera rays themselves
                if (!gen0) {
                    (fastm ? k_raygen<true> : k_raygen<false>)<<<(n_paths + WF_THREADS - 1) / WF_THREADS, WF_THREADS, 0, st>>>(
                        cam, (int)p0, npix, div_npix, n_paths, (uint32_t)(sample_begin + s0), seed_lo, seed_hi, qo[0], qd[0], qb[0], qcount, dyn);
                    ++launches;
                }
                for (int b = 0; b < std::min(max_depth, tail_b); ++b) {
                    const bool prim = gen0 && b == 0;
                    if (timing) spans.push_back({stamp(st), b == 0 ? 2 : 0});
                    if (c->persistent) {
                        uint32_t* work = qcount + (PRT_MAX_DEPTH + 2) + b;
                        (prim ? k_ext0 : k_ext)<<<grid_pt, WF_THREADS, 0, st>>>(sc, qo[cur], qd[cur], hits, qcount, work, b, wide ? c->pt_refill_wide : c->pt_refill, dstats, pg);
                    } else if (c->counters_on) k_extend<true><<<grid, WF_THREADS, 0, st>>>(sc, qo[cur], qd[cur], hits, qcount, b, dstats);
                    else k_extend<false><<<grid, WF_THREADS, 0, st>>>(sc, qo[cur], qd[cur], hits, qcount, b, dstats);
                    if (timing) spans.push_back({stamp(st), 1});
                    auto ksh = prim ? (fastm ? k_shade<true, true> : k_shade<false, true>) : (fastm ? k_shade<true, false> : k_shade<false, false>);
                    ksh<<<grid, WF_THREADS, 0, st>>>(sc, qo[cur], qd[cur], qb[cur], hits, qo[cur ^ 1], qd[cur ^ 1], qb[cur ^ 1],
                                                         contrib, qcount, b, max_depth, pg);
                    if (timing) stamp(st);
                    launches += 2;
                    cur ^= 1;
                    // ray re-sorting (option): the queue this bounce produced, by origin cell and octant
                    if (c->sort_rays && b + 1 < std::min(max_depth, tail_b)) {
                        const unsigned long long h0 = c->hint_rays[0];
                        const double expect = h0 ? (double)n_paths * (double)c->hint_rays[b + 1] / (double)h0 : (double)n_paths;
                        if (expect >= (double)c->sort_min_rays) {
                            if ((rc = sort_ray_queue(c, st, set, n_paths, &qcount[b + 1], qo[cur], qd[cur], qb[cur],
                                                     qo[cur ^ 1], qd[cur ^ 1], qb[cur ^ 1]))) return rc;
                            launches += 2 + 3 * ((std::min(24, std::max(8, c->sort_key_bits)) + 7) / 8);
                            cur ^= 1;
                        }
                    }
                }
            }
            if (tail_b < max_depth) {
                if (timing) spans.push_back({stamp(st), 0});
                const bool tw = wide && !fused;                        // staged scenes with wide nodes: same walk as their extend kernel
                auto kt = fastm ? (c->counters_on ? (brute && fused ? k_tail<true, true, false, true> : tw ? k_tail<true, false, true, true> : k_tail<true, false, false, true>)
                                                  : (brute && fused ? k_tail<false, true, false, true> : tw ? k_tail<false, false, true, true> : k_tail<false, false, false, true>))
                                : (c->counters_on ? (brute && fused ? k_tail<true, true, false, false> : tw ? k_tail<true, false, true, false> : k_tail<true, false, false, false>)
                                                  : (brute && fused ? k_tail<false, true, false, false> : tw ? k_tail<false, false, true, false> : k_tail<false, false, false, false>));
                if (fused && c->tail_pool) {                          // block-local wavefront (scenes of the fused path)
                    auto kp = fastm ? (c->counters_on ? (brute ? k_tail_pool<true, true, true> : k_tail_pool<true, false, true>)
                                                      : (brute ? k_tail_pool<false, true, true> : k_tail_pool<false, false, true>))
                                    : (c->counters_on ? (brute ? k_tail_pool<true, true, false> : k_tail_pool<true, false, false>)
                                                      : (brute ? k_tail_pool<false, true, false> : k_tail_pool<false, false, false>));
                    kp<<<grid, WF_THREADS, 0, st>>>(sc, qo[cur], qd[cur], qb[cur], contrib, qcount, tail_b, max_depth, dstats);
                } else
                kt<<<grid, WF_THREADS, 0, st>>>(sc, qo[cur], qd[cur], qb[cur], contrib, qcount, tail_b, max_depth, dstats);
                if (timing) stamp(st);
                ++launches;
            }
            // resolves read-modify-write the same accumulation pixels: keep them in wave order across the two streams
            if (!direct) {
                if (nsets > 1 && wave > 0) PRT_CUDA(c, cudaStreamWaitEvent(st, c->ev_res[(wave - 1) & 1], 0));
                k_resolve<<<(npix + WF_THREADS - 1) / WF_THREADS, WF_THREADS, 0, st>>>(contrib, (int)p0, npix, S, c->accum);
                if (nsets > 1) PRT_CUDA(c, cudaEventRecord(c->ev_res[wave & 1], st));
                ++launches;
            }
        }
    }
    if (nsets > 1) {
        PRT_CUDA(c, cudaEventRecord(c->ev_join, c->stream2));
        PRT_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    }
    cudaStream_t st = c->stream;
    if (!capturing) PRT_CUDA(c, cudaEventRecord(c->ev1, st));
    // counters come back through pinned memory behind the last kernel: one host wake-up per call, not two
    PRT_CUDA(c, cudaMemcpyAsync(c->stats_host, dstats, (PRT_MAX_DEPTH + 4) * 8, cudaMemcpyDeviceToHost, st));
    if (capturing) {
        cudaGraph_t g = nullptr;
        capturing = false;
        PRT_CUDA(c, cudaStreamEndCapture(st, &g));
        cudaGraphExec_t exec = nullptr;
        cudaError_t e = cudaGraphInstantiate(&exec, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) return set_err(c, PRT_ERR_CUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e));
        if (c->graphs.size() >= 4) { cudaGraphExecDestroy(c->graphs.front().exec); c->graphs.erase(c->graphs.begin()); }   // tail starts flip between neighbours
        c->graphs.push_back(GraphEntry{gkey, exec, launches});
        PRT_CUDA(c, cudaGraphLaunch(exec, st));
        PRT_CUDA(c, cudaEventRecord(c->ev1, st));
    }
    c->pend_launches = launches; c->pend_paths = paths; c->pend_active = true;
    c->accum_spp = (fresh ? 0 : c->accum_spp) + spp;
    if (defer_finish) return PRT_OK;         // multi-device: the caller enqueues the reduce behind this, then finishes every device
    return wavefront_finish(c);
}

// Second half of a render: wait for the stream, read the counters back. Split from the enqueue half so that a multi-device
// context can enqueue every device's waves and the reduce behind them before the host waits for anything.
int wavefront_finish(Ctx* c) {
    if (!c->pend_active) return PRT_OK;
    c->pend_active = false;
    cudaStream_t st = c->stream;
    PRT_CUDA(c, cudaStreamSynchronize(st));
    PRT_CUDA(c, cudaGetLastError());
    const int launches = c->pend_launches; const uint64_t paths = c->pend_paths;
    std::vector<std::pair<size_t, int>>& spans = c->pend_spans;

    const unsigned long long* h = c->stats_host;
    prt_stats& s = c->stats;
    s.rays_total = 0;
    for (int b = 0; b < PRT_MAX_DEPTH; ++b) { s.rays_per_bounce[b] = h[b]; s.rays_total += h[b]; }
    s.nodes_visited = h[PRT_MAX_DEPTH]; s.prims_tested = h[PRT_MAX_DEPTH + 1];
    s.paths = paths;
    s.kernel_launches = launches;
    for (int b = 0; b < PRT_MAX_DEPTH; ++b) c->hint_rays[b] = h[b];
    cudaEventElapsedTime(&s.render_ms, c->ev0, c->ev1);
    float ms[3] = {0.f, 0.f, 0.f};
    int n_primary = 0;
    for (auto& sp : spans) {
        float a = 0.f;
        cudaEventElapsedTime(&a, c->ev_pool[sp.first], c->ev_pool[sp.first + 1]);
        ms[sp.second] += a;
        if (sp.second == 2) ++n_primary;
    }
    // with two waves in flight the per-kernel spans of the two streams overlap in time: their sum can exceed render_ms
    s.extend_ms = ms[0] + ms[2]; s.shade_ms = ms[1]; s.primary_ms = ms[2]; s.primary_launches = n_primary;
    s.n_prims = c->n_prims; s.n_nodes = c->n_prims > 1 ? c->n_prims - 1 : 0;
    return PRT_OK;
}

void wavefront_drop_graphs(Ctx* c) {
    for (GraphEntry& g : c->graphs) cudaGraphExecDestroy(g.exec);
    c->graphs.clear();
    ++c->graph_gen;
}

int wavefront_trace_batch(Ctx* c, const prt_ray* rays, int n, prt_hit* out, int use_bvh) {
    if (n <= 0) return PRT_OK;
    cudaStream_t st = c->stream;
    int rc;
    if ((rc = ensure_wave_buffers(c, std::max<int64_t>(n, 1)))) return rc;
    DevBuf d_rays, d_out;
    if ((rc = dev_reserve(c, d_rays, (size_t)n * sizeof(prt_ray))) || (rc = dev_reserve(c, d_out, (size_t)n * sizeof(prt_hit)))) {
        dev_free(d_rays); dev_free(d_out); return rc;
    }
    SceneView sc = scene_view(c);
    uint32_t* qcount = (uint32_t*)c->qcount.p;
    unsigned long long* dstats = (unsigned long long*)c->stats_dev.p;
    cudaMemcpyAsync(d_rays.p, rays, (size_t)n * sizeof(prt_ray), cudaMemcpyHostToDevice, st);
    cudaMemsetAsync(dstats, 0, (PRT_MAX_DEPTH + 4) * 8, st);
    cudaMemsetAsync(qcount, 0, 2 * (PRT_MAX_DEPTH + 2) * 4, st);
    const int G = (n + WF_THREADS - 1) / WF_THREADS;
    k_load_rays<<<G, WF_THREADS, 0, st>>>((const prt_ray*)d_rays.p, n, (float4*)c->q_o[0].p, (float4*)c->q_d[0].p, qcount);
    const int grid = std::min(G, c->num_sms * c->ctas_per_sm);
    if (use_bvh == 2 && sc.wnodes != nullptr && c->n_prims > 1) {      // the 4-wide quantised nodes
        uint32_t* work = qcount + (PRT_MAX_DEPTH + 2);
        auto k = c->counters_on ? k_extend_wide<true> : k_extend_wide<false>;
        int per_sm = 4;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, WF_THREADS, 0);
        k<<<std::max(1, c->num_sms * std::max(1, per_sm)), WF_THREADS, 0, st>>>(sc, (const float4*)c->q_o[0].p, (const float4*)c->q_d[0].p,
                                                                                (float4*)c->hits.p, qcount, work, 0, c->pt_refill_wide, dstats, PrimaryGen{});
    } else if (use_bvh && c->persistent) {
        uint32_t* work = qcount + (PRT_MAX_DEPTH + 2);
        int per_sm = 4;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_extend_pt<false>, WF_THREADS, 0);
        const int gpt = std::max(1, c->num_sms * std::max(1, per_sm));
        if (c->counters_on)
            k_extend_pt<true><<<gpt, WF_THREADS, 0, st>>>(sc, (const float4*)c->q_o[0].p, (const float4*)c->q_d[0].p, (float4*)c->hits.p, qcount, work, 0, c->pt_refill, dstats, PrimaryGen{});
        else
            k_extend_pt<false><<<gpt, WF_THREADS, 0, st>>>(sc, (const float4*)c->q_o[0].p, (const float4*)c->q_d[0].p, (float4*)c->hits.p, qcount, work, 0, c->pt_refill, dstats, PrimaryGen{});
    } else if (use_bvh) {
        if (c->counters_on)
            k_extend<true><<<grid, WF_THREADS, 0, st>>>(sc, (const float4*)c->q_o[0].p, (const float4*)c->q_d[0].p, (float4*)c->hits.p, qcount, 0, dstats);
        else
            k_extend<false><<<grid, WF_THREADS, 0, st>>>(sc, (const float4*)c->q_o[0].p, (const float4*)c->q_d[0].p, (float4*)c->hits.p, qcount, 0, dstats);
    } else {
        k_extend_brute<<<grid, WF_THREADS, 0, st>>>(sc, (const uint32_t*)c->vals[c->sorted_slot].p, (const float4*)c->q_o[0].p,
                                                    (const float4*)c->q_d[0].p, (float4*)c->hits.p, qcount, 0);
    }
    k_hit_records<<<G, WF_THREADS, 0, st>>>(sc, (const float4*)c->q_o[0].p, (const float4*)c->q_d[0].p,
                                            (const float4*)c->hits.p, n, (prt_hit*)d_out.p);
    cudaMemcpyAsync(out, d_out.p, (size_t)n * sizeof(prt_hit), cudaMemcpyDeviceToHost, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    unsigned long long h[PRT_MAX_DEPTH + 4];
    if (e == cudaSuccess) e = cudaMemcpy(h, dstats, sizeof(h), cudaMemcpyDeviceToHost);
    dev_free(d_rays); dev_free(d_out);
    if (e != cudaSuccess) return set_err(c, PRT_ERR_CUDA, std::string("trace_batch: ") + cudaGetErrorString(e));
    c->stats.nodes_visited = h[PRT_MAX_DEPTH]; c->stats.prims_tested = h[PRT_MAX_DEPTH + 1];
    c->stats.rays_total = h[0];
    return PRT_OK;
}

// ------------------------------------------------------------------ material known answers (parity entry)
// MaterialHandle::{Emit,Scatter} (material.h:139-161) for n independent records through the SAME device function the
// render path calls (prt_device.cuh:scatter): the reference's 600 known answers (tests/golden) are checked against it.
__global__ void k_debug_scatter(const prt_material* __restrict__ mats, const float* __restrict__ in_dirs, const float* __restrict__ normals,
                                const int* __restrict__ fronts, const uint32_t* __restrict__ keys, uint32_t seed_lo, uint32_t seed_hi,
                                int n, float* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const prt_material pm = mats[i];
    MatRec m; m.a = make_float4(pm.r, pm.g, pm.b, pm.param); m.b = make_float4(__int_as_float(pm.type), 0.f, 0.f, 0.f);
    Surface sf; sf.p = mk(0.f, 0.f, 0.f); sf.n = mk(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]); sf.front = fronts[i] != 0; sf.mat = 0;
    Rng rng = rng_for_bounce(path_key(seed_lo, seed_hi, keys[3 * i], keys[3 * i + 1]), keys[3 * i + 2]);
    V3 atten = mk(0.f, 0.f, 0.f), od = mk(0.f, 0.f, 0.f), emit;
    const bool ok = scatter(m, mk(in_dirs[3 * i], in_dirs[3 * i + 1], in_dirs[3 * i + 2]), sf, rng, atten, od, emit);
    if (!ok) od = mk(0.f, 0.f, 0.f);
    float* o = out + 10 * (size_t)i;
    o[0] = ok ? 1.f : 0.f; o[1] = atten.x; o[2] = atten.y; o[3] = atten.z; o[4] = od.x; o[5] = od.y; o[6] = od.z;
    o[7] = emit.x; o[8] = emit.y; o[9] = emit.z;
}

int wavefront_debug_scatter(Ctx* c, const prt_material* mats, const float* in_dirs, const float* normals, const int32_t* fronts,
                            const uint32_t* keys, uint64_t seed, int n, float* out) {
    if (n <= 0) return PRT_OK;
    DevBuf dm, di, dn, df, dk, dout;
    int rc = 0;
    const size_t N = (size_t)n;
    if ((rc = dev_reserve(c, dm, N * sizeof(prt_material))) || (rc = dev_reserve(c, di, N * 12)) || (rc = dev_reserve(c, dn, N * 12)) ||
        (rc = dev_reserve(c, df, N * 4)) || (rc = dev_reserve(c, dk, N * 12)) || (rc = dev_reserve(c, dout, N * 40))) {
        for (DevBuf* b : {&dm, &di, &dn, &df, &dk, &dout}) dev_free(*b);
        return rc;
    }
    cudaStream_t st = c->stream;
    cudaMemcpyAsync(dm.p, mats, N * sizeof(prt_material), cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(di.p, in_dirs, N * 12, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(dn.p, normals, N * 12, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(df.p, fronts, N * 4, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(dk.p, keys, N * 12, cudaMemcpyHostToDevice, st);
    k_debug_scatter<<<(n + 127) / 128, 128, 0, st>>>((const prt_material*)dm.p, (const float*)di.p, (const float*)dn.p, (const int*)df.p,
                                                     (const uint32_t*)dk.p, (uint32_t)seed, (uint32_t)(seed >> 32), n, (float*)dout.p);
    cudaMemcpyAsync(out, dout.p, N * 40, cudaMemcpyDeviceToHost, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    for (DevBuf* b : {&dm, &di, &dn, &df, &dk, &dout}) dev_free(*b);
    if (e != cudaSuccess) return set_err(c, PRT_ERR_CUDA, std::string("debug_scatter: ") + cudaGetErrorString(e));
    return PRT_OK;
}

void wavefront_set_stack_limit(int limit) {
#if PRT_CHECKED
    limit = limit < STACK_SMEM ? STACK_SMEM : (limit > STACK_TOTAL ? STACK_TOTAL : limit);
    cudaMemcpyToSymbol(g_prt_stack_limit, &limit, sizeof limit);
#else
    (void)limit;
#endif
}

unsigned wavefront_check_flags() {
#if PRT_CHECKED
    unsigned v = 0, z = 0;
    cudaMemcpyFromSymbol(&v, g_prt_check_flags, sizeof v);
    cudaMemcpyToSymbol(g_prt_check_flags, &z, sizeof z);
    return v;
#else
    return 0u;
#endif
}

__global__ void k_scale(float* a, size_t n, float s) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] *= s;
}
int film_scale_accum(Ctx* c, float s) {
    size_t n = (size_t)c->cam.width * c->cam.height * 3;
    k_scale<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->accum, n, s);
    PRT_CUDA(c, cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

int film_resolve_display(Ctx* c, float exposure, float gamma, uint8_t* host_rgba) {
    const int npix = c->cam.width * c->cam.height;
    int rc;
    if ((rc = dev_reserve(c, c->display, (size_t)npix * 4))) return rc;
    float w = (float)c->accum_spp;
    k_display<<<(npix + 255) / 256, 256, 0, c->stream>>>(c->accum, npix, w > 0.f ? 1.0f / w : 0.f, w, exposure,
                                                        1.0f / gamma, (uint8_t*)c->display.p);
    PRT_CUDA(c, cudaMemcpyAsync(host_rgba, c->display.p, (size_t)npix * 4, cudaMemcpyDeviceToHost, c->stream));
    PRT_CUDA(c, cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

}  // namespace prt

// ------------------------------------------------------------------ closest-hit batch without any allocation
// prt_trace_batch_static (prt.h): the kernel chain of wavefront_trace_batch (k_load_rays -> extend -> k_hit_records, the same
// kernels, hence the same bits) with its scratch - the SoA ray queue, the hit plane and the queue / work counters - in a
// file-scope device array instead of context buffers that grow through cudaMalloc. The array exists once per GPU that
// loads the library (3 MB + 1 KB, held until the process exits), every call on that GPU shares it and it keeps the last
// call's contents: the counters are cleared by every call, the planes are fully overwritten for [0, n) before they are read.
namespace prt {
constexpr int    NOALLOC_MAX_RAYS  = PRT_TRACE_STATIC_MAX_RAYS;
constexpr size_t NOALLOC_CTR_BYTES = 1024;                                       // qcount + work (272 B) | counters (288 B at +512)
constexpr size_t NOALLOC_BYTES     = NOALLOC_CTR_BYTES + (size_t)NOALLOC_MAX_RAYS * 3 * sizeof(float4);
static_assert(2 * (PRT_MAX_DEPTH + 2) * 4 <= 512 && (PRT_MAX_DEPTH + 4) * 8 <= 512, "counter block of the static scratch");
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

// In host code `g_noalloc_scratch` is only a handle of the symbol, not a device address: the address each device maps it at
// comes from cudaGetSymbolAddress with that device current, once per device.
constexpr int NOALLOC_MAX_DEVICES = 64;
static unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mu;

static int noalloc_scratch(Ctx* c, unsigned char** out) {
    if (c->device < 0 || c->device >= NOALLOC_MAX_DEVICES) return set_err(c, PRT_ERR_INVALID, "prt_trace_batch_static: device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[c->device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);             // c->device is current (set by the caller)
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return set_err(c, PRT_ERR_CUDA, std::string("prt_trace_batch_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        }
        g_noalloc_ptr[c->device] = (unsigned char*)p;
    }
    *out = g_noalloc_ptr[c->device];
    return PRT_OK;
}

static int trace_batch_static(Ctx* c, const prt_ray* d_rays, int n, prt_hit* d_out, int use_bvh, cudaStream_t st) {
    unsigned char* base = nullptr;
    int rc;
    if ((rc = noalloc_scratch(c, &base))) return rc;
    if (n == 0) return PRT_OK;                                                   // set-up only
    uint32_t* qcount = (uint32_t*)base;
    uint32_t* work = qcount + (PRT_MAX_DEPTH + 2);
    unsigned long long* dstats = (unsigned long long*)(base + 512);
    float4* qo = (float4*)(base + NOALLOC_CTR_BYTES);
    float4* qd = qo + NOALLOC_MAX_RAYS;
    float4* hits = qd + NOALLOC_MAX_RAYS;
    SceneView sc = scene_view(c);
    PRT_CUDA(c, cudaMemsetAsync(base, 0, NOALLOC_CTR_BYTES, st));                // counters must start at zero
    const int G = (n + WF_THREADS - 1) / WF_THREADS;
    k_load_rays<<<G, WF_THREADS, 0, st>>>(d_rays, n, qo, qd, qcount);
    const int grid = std::min(G, c->num_sms * c->ctas_per_sm);
    if (use_bvh == 2 && sc.wnodes != nullptr && c->n_prims > 1) {
        int per_sm = 4;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_extend_wide<false>, WF_THREADS, 0);
        k_extend_wide<false><<<std::max(1, c->num_sms * std::max(1, per_sm)), WF_THREADS, 0, st>>>(
            sc, qo, qd, hits, qcount, work, 0, c->pt_refill_wide, dstats, PrimaryGen{});
    } else if (use_bvh && c->persistent) {
        int per_sm = 4;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_extend_pt<false>, WF_THREADS, 0);
        k_extend_pt<false><<<std::max(1, c->num_sms * std::max(1, per_sm)), WF_THREADS, 0, st>>>(
            sc, qo, qd, hits, qcount, work, 0, c->pt_refill, dstats, PrimaryGen{});
    } else if (use_bvh) {
        k_extend<false><<<grid, WF_THREADS, 0, st>>>(sc, qo, qd, hits, qcount, 0, dstats);
    } else {
        k_extend_brute<<<grid, WF_THREADS, 0, st>>>(sc, (const uint32_t*)c->vals[c->sorted_slot].p, qo, qd, hits, qcount, 0);
    }
    k_hit_records<<<G, WF_THREADS, 0, st>>>(sc, qo, qd, hits, n, d_out);
    PRT_CUDA(c, cudaGetLastError());
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    PRT_CUDA(c, cudaStreamIsCapturing(st, &cap));
    if (cap == cudaStreamCaptureStatusNone) PRT_CUDA(c, cudaStreamSynchronize(st));   // a capturing stream only records the chain
    return PRT_OK;
}
}  // namespace prt

extern "C" int prt_trace_batch_static(prt_ctx* ctx, const prt_ray* rays_dev, int n, prt_hit* out_dev, int use_bvh, void* cuda_stream) {
    using namespace prt;
    if (!ctx || (n > 0 && (!rays_dev || !out_dev))) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_trace_batch_static: NULL argument");
    Ctx* c = &ctx->c;
    if (n < 0) return set_err(c, PRT_ERR_INVALID, "prt_trace_batch_static: negative ray count");
    if (n > NOALLOC_MAX_RAYS) {
        char buf[160];
        snprintf(buf, sizeof buf, "prt_trace_batch_static: %d rays exceed the static scratch (PRT_TRACE_STATIC_MAX_RAYS = %d); split the batch or use prt_trace_batch",
                 n, NOALLOC_MAX_RAYS);
        return set_err(c, PRT_ERR_INVALID, buf);
    }
    if (use_bvh < 0 || use_bvh > 2) return set_err(c, PRT_ERR_INVALID, "prt_trace_batch_static: use_bvh must be 0, 1 or 2");
    if (!c->scene_set || !c->bvh_built) return set_err(c, PRT_ERR_INVALID, "prt_trace_batch_static: scene/BVH not ready");
    PRT_CUDA(c, cudaSetDevice(c->device));
    return trace_batch_static(c, rays_dev, n, out_dev, use_bvh, cuda_stream ? (cudaStream_t)cuda_stream : c->stream);
}
