// lbvh.cu — GPU LBVH build: primitive AABBs -> scene bounds -> 30-bit Morton codes ->
// stable LSD radix sort (8-bit digits) -> Karras 2012 hierarchy -> bottom-up refit ->
// 64-byte two-child traversal nodes + leaf-slot-ordered triangle planes.
//
// The reference has no acceleration structure: PrimitiveList::Intersect loops over every
// primitive (src/core/primitive.cpp:26). This build is what makes that loop O(log n).
// Integer outputs (Morton codes, sorted order, topology) are bit-exact against the CPU
// restatement in oracle/prt_oracle.cpp (tests/test_gpu_parity.py::test_lbvh_bit_exact).
#include "prt_internal.h"
#include <cooperative_groups.h>
#include <algorithm>

namespace cg = cooperative_groups;

namespace prt {

// ------------------------------------------------------------------ helpers
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// b[0..5] = bounds of the primitives' boxes (scene bounds), b[6..11] = bounds of the box CENTRES (Morton quantisation grid)
__global__ void k_init_bounds(int* b) {
    if (threadIdx.x < 12) b[threadIdx.x] = (threadIdx.x % 6) < 3 ? 0x7fffffff : (int)0x80000000;   // min = +max ordered, max = lowest
}

// Block reduction of the scene bounds (bounds[0..5]) and of the bounds of the box centres (bounds[6..11]) over the boxes of
// one block's threads, then one ordered-int atomic per value and block (min/max are exact, so the order does not matter).
__device__ __forceinline__ void reduce_bounds(const float (&lo)[3], const float (&hi)[3], bool valid, int* __restrict__ bounds) {
    __shared__ float s_lo[6][8], s_hi[6][8];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        float a = lo[k % 3], b = hi[k % 3];
        if (k >= 3) { a = valid ? (lo[k - 3] + hi[k - 3]) * 0.5f : a; b = valid ? a : b; }      // centre, as k_morton computes it
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a = fminf(a, __shfl_xor_sync(0xffffffffu, a, o));
            b = fmaxf(b, __shfl_xor_sync(0xffffffffu, b, o));
        }
        if (lane == 0) { s_lo[k][w] = a; s_hi[k][w] = b; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        int k = threadIdx.x;
        float a = s_lo[k][0], b = s_hi[k][0];
        for (int i = 1; i < (int)(blockDim.x >> 5); ++i) { a = fminf(a, s_lo[k][i]); b = fmaxf(b, s_hi[k][i]); }
        const int at = k < 3 ? k : 6 + (k - 3);
        if (a <= b) { atomicMin(&bounds[at], f2ord(a)); atomicMax(&bounds[at + 3], f2ord(b)); }
    }
}

// AABB of primitive q (input order: spheres, quads, triangles). The exact expression
// order is part of the bit-exact contract with the oracle.
// Triangles: their boxes (and their share of the two bounds, tri_bounds) were written when the scene was ingested
// (k_ingest_tris had the vertices in registers), so this pass runs over the spheres and quads only.
__global__ void k_prim_bounds(const SphereRec* __restrict__ spheres, int ns,
                              const QuadRec* __restrict__ quads, int nq,
                              const float4* __restrict__ tri_v, int nt, const int* __restrict__ tri_bounds,
                              float4* __restrict__ amin, float4* __restrict__ amax,
                              int* __restrict__ bounds) {
    if (tri_bounds && blockIdx.x == 0 && threadIdx.x < 12) {
        const int k = threadIdx.x, v = tri_bounds[k];
        if ((k % 6) < 3) atomicMin(&bounds[k], v); else atomicMax(&bounds[k], v);
    }
    int n = ns + nq + nt;
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {__int_as_float(0x7f800000), __int_as_float(0x7f800000), __int_as_float(0x7f800000)};
    float hi[3] = {__int_as_float(0xff800000), __int_as_float(0xff800000), __int_as_float(0xff800000)};
    if (q < n) {
        if (q < ns) {
            float4 a = spheres[q].a;
            lo[0] = a.x - a.w; lo[1] = a.y - a.w; lo[2] = a.z - a.w;
            hi[0] = a.x + a.w; hi[1] = a.y + a.w; hi[2] = a.z + a.w;
        } else if (q < ns + nq) {
            QuadRec r = quads[q - ns];
            float c[3] = {r.a.x, r.a.y, r.a.z};
            float a[3] = {r.b.x * r.a.w, r.b.y * r.a.w, r.b.z * r.a.w};   // u * half_w
            float b[3] = {r.c.x * r.b.w, r.c.y * r.b.w, r.c.z * r.b.w};   // v * half_h
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                float p0 = (c[k] + a[k]) + b[k], p1 = (c[k] + a[k]) - b[k];
                float p2 = (c[k] - a[k]) + b[k], p3 = (c[k] - a[k]) - b[k];
                lo[k] = fminf(fminf(p0, p1), fminf(p2, p3));
                hi[k] = fmaxf(fmaxf(p0, p1), fmaxf(p2, p3));
            }
        } else {                                                         // only without an ingest pass (never on the product path)
            int t = q - ns - nq;
            float4 v0 = tri_v[6 * (size_t)t], v1 = tri_v[6 * (size_t)t + 1], v2 = tri_v[6 * (size_t)t + 2];
            lo[0] = fminf(fminf(v0.x, v1.x), v2.x); hi[0] = fmaxf(fmaxf(v0.x, v1.x), v2.x);
            lo[1] = fminf(fminf(v0.y, v1.y), v2.y); hi[1] = fmaxf(fmaxf(v0.y, v1.y), v2.y);
            lo[2] = fminf(fminf(v0.z, v1.z), v2.z); hi[2] = fmaxf(fmaxf(v0.z, v1.z), v2.z);
        }
        amin[2 * q] = make_float4(lo[0], lo[1], lo[2], 0.f);            // {min, max} interleaved: one 32-byte sector per box
        amax[2 * q] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
    reduce_bounds(lo, hi, q < n, bounds);
}

// Triangle ingest (prt_set_scene): the caller's packed arrays (9 floats per triangle, optional 9 floats of vertex normals,
// optional material index) are copied to the device as they are and expanded here into the 96-byte records the build
// gathers {v0, v1 | material, v2, n0, n1, n2}. The vertices are in registers, so the triangle's box (build input, same
// min/max as the oracle's prim_bounds) and its share of the scene / centre bounds are produced here as well: the build has
// no pass over the triangles before the Morton codes. Flat normal when none is given: cross(v1 - v0, v2 - v0), unnormalised.
__global__ void __launch_bounds__(256)
k_ingest_tris(int n, const float* __restrict__ v, const float* __restrict__ nrm, const int* __restrict__ mat, int mat_default,
              float4* __restrict__ out_v, float4* __restrict__ out_n, float4* __restrict__ box, int* __restrict__ tri_bounds, int positions_only) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {__int_as_float(0x7f800000), __int_as_float(0x7f800000), __int_as_float(0x7f800000)};
    float hi[3] = {__int_as_float(0xff800000), __int_as_float(0xff800000), __int_as_float(0xff800000)};
    if (t < n) {
        float a[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) a[k] = v[9 * (size_t)t + k];
#pragma unroll
        for (int k = 0; k < 3; ++k)
            out_v[6 * (size_t)t + k] = make_float4(a[3 * k], a[3 * k + 1], a[3 * k + 2], k == 1 ? __int_as_float(mat ? mat[t] : mat_default) : 0.f);
        if (positions_only) {
            // prt_set_scene_build: normals and materials are still on their way over PCIe; k_finish_tris adds them behind the build
        } else if (nrm) {
#pragma unroll
            for (int k = 0; k < 3; ++k)
                out_n[6 * (size_t)t + k] = make_float4(nrm[9 * (size_t)t + 3 * k], nrm[9 * (size_t)t + 3 * k + 1], nrm[9 * (size_t)t + 3 * k + 2], 0.f);
        } else {
            float e1x = a[3] - a[0], e1y = a[4] - a[1], e1z = a[5] - a[2], e2x = a[6] - a[0], e2y = a[7] - a[1], e2z = a[8] - a[2];
            float nx = e1y * e2z - e2y * e1z, ny = e1z * e2x - e2z * e1x, nz = e1x * e2y - e2x * e1y;
#pragma unroll
            for (int k = 0; k < 3; ++k) out_n[6 * (size_t)t + k] = make_float4(nx, ny, nz, 0.f);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) { lo[k] = fminf(fminf(a[k], a[3 + k]), a[6 + k]); hi[k] = fmaxf(fmaxf(a[k], a[3 + k]), a[6 + k]); }
        box[2 * (size_t)t] = make_float4(lo[0], lo[1], lo[2], 0.f);
        box[2 * (size_t)t + 1] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
    reduce_bounds(lo, hi, t < n, tri_bounds);
}

// raw_* = the packed arrays already on the device; allocates the records and the primitive boxes (all primitives)
int lbvh_ingest_tris(Ctx* c, int n_spheres, int n_quads, int n_tris, bool have_normals, bool have_mat, int mat_default, bool positions_only) {
    int rc;
    const size_t nn = (size_t)n_spheres + n_quads + n_tris, nt = (size_t)n_tris;
    if ((rc = dev_reserve(c, c->tri_in_v, nt * 96)) || (rc = dev_reserve(c, c->aabb_min, nn * 32)) || (rc = dev_reserve(c, c->tri_bounds, 12 * sizeof(int))))
        return rc;
    k_init_bounds<<<1, 32, 0, c->stream>>>((int*)c->tri_bounds.p);
    if (n_tris > 0)
        k_ingest_tris<<<(unsigned)((nt + 255) / 256), 256, 0, c->stream>>>(n_tris, (const float*)c->raw_v.p, have_normals ? (const float*)c->raw_n.p : nullptr,
            have_mat && !positions_only ? (const int*)c->raw_m.p : nullptr, mat_default, (float4*)c->tri_in_v.p, (float4*)c->tri_in_v.p + 3,
            (float4*)c->aabb_min.p + 2 * ((size_t)n_spheres + n_quads), (int*)c->tri_bounds.p, positions_only ? 1 : 0);
    return PRT_OK;
}

__device__ __forceinline__ uint32_t expand10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__device__ __forceinline__ uint32_t quant10(float c, float lo, float hi) {
    float e = hi - lo;
    float x = e > 0.0f ? (c - lo) / e : 0.0f;
    int xi = (int)(x * 1024.0f);
    xi = xi < 0 ? 0 : (xi > 1023 ? 1023 : xi);
    return (uint32_t)xi;
}

__global__ void k_morton(const float4* __restrict__ amin, const float4* __restrict__ amax, int n,
                         const int* __restrict__ bounds, uint32_t* __restrict__ keys,
                         uint32_t* __restrict__ vals, uint32_t* __restrict__ morton_out) {
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    float lx = ord2f(bounds[0]), ly = ord2f(bounds[1]), lz = ord2f(bounds[2]);
    float hx = ord2f(bounds[3]), hy = ord2f(bounds[4]), hz = ord2f(bounds[5]);
    float4 a = amin[2 * q], b = amax[2 * q];
    float cx = (a.x + b.x) * 0.5f, cy = (a.y + b.y) * 0.5f, cz = (a.z + b.z) * 0.5f;
    uint32_t code = (expand10(quant10(cx, lx, hx)) << 2) | (expand10(quant10(cy, ly, hy)) << 1) |
                    expand10(quant10(cz, lz, hz));
    keys[q] = code; vals[q] = (uint32_t)q; morton_out[q] = code;
}

// ------------------------------------------------------------------ radix sort (stable LSD)
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;   // 4096 keys per block
constexpr int RS_WARPS = RS_THREADS / 32;

__global__ void __launch_bounds__(RS_THREADS)
k_rs_hist(const uint32_t* __restrict__ keys, int n, int shift, uint32_t* __restrict__ block_hist, int nblocks,
          const uint32_t* __restrict__ n_dev = nullptr) {
    __shared__ uint32_t h[256];
    if (n_dev) {                       // ray queues: the length lives on the device, the grid is sized for the capacity
        n = (int)*n_dev; nblocks = (n + RS_TILE - 1) / RS_TILE;
        if ((int)blockIdx.x >= nblocks) return;
    }
    h[threadIdx.x] = 0;
    __syncthreads();
    int base = blockIdx.x * RS_TILE;
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        int i = base + r * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    block_hist[(size_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// Row scan: block_hist is digit-major [256][nblocks]. CTA d turns row d into exclusive prefixes over
// the blocks (in place) and writes the row total to digit_total[d]; the 256 digit bases are then a
// 256-element scan that every scatter CTA redoes in shared memory. (A single-CTA scan of all
// 256*nblocks counters took 574 us per pass at 10 M keys: 40 % of the whole build.)
__global__ void __launch_bounds__(256) k_rs_scan_rows(uint32_t* __restrict__ hist, int nblocks, uint32_t* __restrict__ digit_total,
                                                      const uint32_t* __restrict__ n_dev = nullptr) {
    __shared__ uint32_t warp_sum[8];
    if (n_dev) nblocks = ((int)*n_dev + RS_TILE - 1) / RS_TILE;
    uint32_t* row = hist + (size_t)blockIdx.x * nblocks;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int per = (nblocks + 255) / 256;
    const int lo = min(tid * per, nblocks), hi = min(lo + per, nblocks);
    uint32_t s = 0;
    for (int i = lo; i < hi; ++i) s += row[i];
    uint32_t inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) warp_sum[w] = inc;
    __syncthreads();
    uint32_t wbase = 0, total = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) { uint32_t v = warp_sum[k]; if (k < w) wbase += v; total += v; }
    uint32_t run = wbase + (inc - s);
    for (int i = lo; i < hi; ++i) { uint32_t v = row[i]; row[i] = run; run += v; }
    if (tid == 0) digit_total[blockIdx.x] = total;
}

// lanes of the warp that hold the same 9-bit digit (8 digit bits + "invalid"): nine ballots, constant cost; PRT_RS_MATCH_HW=1
// selects the MATCH.ANY instruction instead (its cost grows with the number of distinct values: ~30 for random digits)
__device__ __forceinline__ uint32_t rs_match(uint32_t digit) {
#if PRT_RS_MATCH_HW
    return __match_any_sync(0xffffffffu, digit);
#else
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 9; ++b) {
        const bool bit = (digit >> b) & 1u;
        const uint32_t bal = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
#endif
}

// Scatter pass. Each warp ranks its 512 keys with __match_any_sync (stable within the warp's contiguous chunk); the tile is
// then put in digit order IN SHARED MEMORY and written out run by run, so that consecutive threads store to consecutive
// addresses (round 1 stored straight from registers: consecutive lanes hit different bins, 4-byte scattered stores, 4.2 GB
// of L2-level traffic for 0.55 GB of DRAM and 780 GB/s).
__global__ void __launch_bounds__(RS_THREADS)
k_rs_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
             uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n, int shift,
             const uint32_t* __restrict__ block_off, int nblocks, const uint32_t* __restrict__ digit_total,
             const uint32_t* __restrict__ n_dev = nullptr) {
    __shared__ uint32_t cnt[RS_WARPS][256];
    if (n_dev) {
        n = (int)*n_dev; nblocks = (n + RS_TILE - 1) / RS_TILE;
        if ((int)blockIdx.x >= nblocks) return;
    }
    __shared__ uint32_t gbase[256], lbase[256];
    __shared__ uint32_t dwarp[8];
    __shared__ uint32_t s_key[RS_TILE], s_val[RS_TILE];
    int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int i = tid; i < RS_WARPS * 256; i += RS_THREADS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    uint32_t key[RS_ITEMS], val[RS_ITEMS], rank[RS_ITEMS];   // (fetching the values only when their slot is known saves 21 registers and costs 18 %: 570 vs 484 us)
    const int tile0 = blockIdx.x * RS_TILE, base = tile0 + w * (RS_ITEMS * 32);
    const int tile_n = min(RS_TILE, n - tile0);
    uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        int i = base + r * 32 + lane;
        bool valid = i < n;
        key[r] = valid ? keys_in[i] : 0u;
        val[r] = valid ? vals_in[i] : 0u;
        uint32_t digit = valid ? ((key[r] >> shift) & 255u) : 256u;
        uint32_t peers = rs_match(digit);
        int leader = __ffs(peers) - 1;
        uint32_t b = 0;
        if (valid && lane == leader) { b = cnt[w][digit]; cnt[w][digit] = b + __popc(peers); }
        b = __shfl_sync(0xffffffffu, b, leader);
        rank[r] = b + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();
    {   // per digit (one thread each): count in this tile -> exclusive scan over the digits = position of the digit's run in the
        // tile; global start of that run = digit base (scan of the 256 totals) + this block's offset within the digit
        uint32_t tile_cnt = 0;
#pragma unroll
        for (int ww = 0; ww < RS_WARPS; ++ww) tile_cnt += cnt[ww][tid];
        uint32_t v = digit_total[tid], inc = v, linc = tile_cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, o), u = __shfl_up_sync(0xffffffffu, linc, o);
            if (lane >= o) { inc += t; linc += u; }
        }
        if (lane == 31) { dwarp[w] = inc; gbase[w] = linc; }            // gbase[0..7] borrowed for the warp sums of the tile counts
        __syncthreads();
        uint32_t wb = 0, lwb = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) if (k < w) { wb += dwarp[k]; lwb += gbase[k]; }
        __syncthreads();
        const uint32_t lb = lwb + linc - tile_cnt;
        lbase[tid] = lb;
        gbase[tid] = (wb + inc - v) + block_off[(size_t)tid * nblocks + blockIdx.x];
        uint32_t run = lb;                                              // tile-local start of every warp's share of the digit
#pragma unroll
        for (int ww = 0; ww < RS_WARPS; ++ww) { uint32_t c = cnt[ww][tid]; cnt[ww][tid] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        int i = base + r * 32 + lane;
        if (i < n) {
            const uint32_t lp = cnt[w][(key[r] >> shift) & 255u] + rank[r];
            if (!PRT_CHECK(lp < (uint32_t)RS_TILE, PRT_CHK_SORT)) continue;
            s_key[lp] = key[r]; s_val[lp] = val[r];
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        const int idx = r * RS_THREADS + tid;
        if (idx < tile_n) {
            const uint32_t k = s_key[idx], d = (k >> shift) & 255u;
            const uint32_t pos = gbase[d] + ((uint32_t)idx - lbase[d]);
            if (!PRT_CHECK(pos < (uint32_t)n, PRT_CHK_SORT)) continue;
            keys_out[pos] = k; vals_out[pos] = s_val[idx];
        }
    }
}

// ------------------------------------------------------------------ ray re-sorting (the reference lists it as wanted, wavefront.md:86-96)
// After a diffuse bounce the queue order is the order in which paths happened to finish shading: neighbouring lanes start
// anywhere in the scene. Sorting the queue by the Morton code of the ORIGIN cell (then the direction octant) puts rays
// that start in the same place and head the same way into the same warp. The order of a queue has no influence on any
// path (RNG is keyed by pixel/sample/bounce, contributions are slot-indexed): the image is bit-identical with and without.
__global__ void k_ray_keys(const float4* __restrict__ qo, const float4* __restrict__ qd, const uint32_t* __restrict__ n_dev,
                           const int* __restrict__ bounds, int key_bits, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const uint32_t n = *n_dev;
    const float lx = ord2f(bounds[0]), ly = ord2f(bounds[1]), lz = ord2f(bounds[2]);
    const float hx = ord2f(bounds[3]), hy = ord2f(bounds[4]), hz = ord2f(bounds[5]);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 o = __ldcs(&qo[i]), d = __ldcs(&qd[i]);
        const uint32_t m = (expand10(quant10(o.x, lx, hx)) << 2) | (expand10(quant10(o.y, ly, hy)) << 1) | expand10(quant10(o.z, lz, hz));
        const uint32_t oct = (d.x < 0.f ? 4u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 1u : 0u);
        keys[i] = ((m >> (33 - key_bits)) << 3) | oct;         // key_bits - 3 Morton bits above the octant
        vals[i] = i;
    }
}

__global__ void k_ray_gather(const uint32_t* __restrict__ order, const uint32_t* __restrict__ n_dev,
                             const float4* __restrict__ so, const float4* __restrict__ sd, const float4* __restrict__ sb,
                             float4* __restrict__ to, float4* __restrict__ td, float4* __restrict__ tb) {
    const uint32_t n = *n_dev;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t j = order[i];
        const float4 o = __ldcs(&so[j]), d = __ldcs(&sd[j]), b = __ldcs(&sb[j]);
        __stcs(&to[i], o); __stcs(&td[i], d); __stcs(&tb[i], b);
    }
}

// Sort the ray queue (so, sd, sb)[0 .. *n_dev) into (to, td, tb). `capacity` sizes the grids; nothing returns to the host.
int sort_ray_queue(Ctx* c, cudaStream_t st, int set, int64_t capacity, const uint32_t* n_dev, const float4* so, const float4* sd,
                   const float4* sb, float4* to, float4* td, float4* tb) {
    int rc;
    const size_t cap = (size_t)capacity;
    DevBuf* rk = c->ray_keys[set]; DevBuf* rv = c->ray_vals[set];
    const int nblocks = (int)((cap + RS_TILE - 1) / RS_TILE);
    for (int k = 0; k < 2; ++k)
        if ((rc = dev_reserve(c, rk[k], cap * 4)) || (rc = dev_reserve(c, rv[k], cap * 4))) return rc;
    if ((rc = dev_reserve(c, c->ray_hist[set], ((size_t)nblocks * 256 + 256) * 4))) return rc;
    const int key_bits = std::min(24, std::max(8, c->sort_key_bits));
    const int passes = (key_bits + 7) / 8;
    const int G = (int)std::min<size_t>((cap + 255) / 256, (size_t)c->num_sms * 16);
    k_ray_keys<<<G, 256, 0, st>>>(so, sd, n_dev, (const int*)c->scene_bounds.p, key_bits, (uint32_t*)rk[0].p, (uint32_t*)rv[0].p);
    uint32_t* hist = (uint32_t*)c->ray_hist[set].p;
    int cur = 0;
    for (int pass = 0; pass < passes; ++pass) {
        const int shift = pass * 8;
        k_rs_hist<<<nblocks, RS_THREADS, 0, st>>>((const uint32_t*)rk[cur].p, 0, shift, hist, nblocks, n_dev);
        k_rs_scan_rows<<<256, 256, 0, st>>>(hist, nblocks, hist + (size_t)nblocks * 256, n_dev);
        k_rs_scatter<<<nblocks, RS_THREADS, 0, st>>>((const uint32_t*)rk[cur].p, (const uint32_t*)rv[cur].p, (uint32_t*)rk[cur ^ 1].p,
                                                     (uint32_t*)rv[cur ^ 1].p, 0, shift, hist, nblocks, hist + (size_t)nblocks * 256, n_dev);
        cur ^= 1;
    }
    k_ray_gather<<<G, 256, 0, st>>>((const uint32_t*)rv[cur].p, n_dev, so, sd, sb, to, td, tb);
    return PRT_OK;
}

// ------------------------------------------------------------------ Karras hierarchy
__device__ __forceinline__ int delta(const uint32_t* __restrict__ k, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint32_t a = k[i], b = k[j];
    if (a == b) return 32 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clz(a ^ b);
}

// children[2i], children[2i+1]: >= 0 internal node, < 0 = ~leaf position
__global__ void k_karras(const uint32_t* __restrict__ keys, int n, int* __restrict__ children,
                         int* __restrict__ parent, int* __restrict__ leaf_parent) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    if (i == 0) parent[0] = -1;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    int left, right;
    if (lo == gamma) { left = ~gamma; leaf_parent[gamma] = i; } else { left = gamma; parent[gamma] = i; }
    if (hi == gamma + 1) { right = ~(gamma + 1); leaf_parent[gamma + 1] = i; } else { right = gamma + 1; parent[gamma + 1] = i; }
    children[2 * i] = left; children[2 * i + 1] = right;
}

__device__ __forceinline__ int leaf_code(int pos, uint32_t q, int ns, int nq) {
    uint32_t payload = q < (uint32_t)ns ? PRT_MAKE_LEAF(0, q)
                     : q < (uint32_t)(ns + nq) ? PRT_MAKE_LEAF(1, q - ns)
                     : PRT_MAKE_LEAF(2, pos);
    return ~(int)payload;
}

// bottom-up refit: the second thread to reach a node merges the child boxes and, having both child boxes
// in registers, writes the 64-byte traversal node as well (no separate emit pass)
// amin / amax: the leaf boxes in sorted (leaf-slot) order, {min, max} pairs written by k_gather_tris
__global__ void k_refit(int n, const uint32_t* __restrict__ sorted_prim, const float4* __restrict__ amin,
                        const float4* __restrict__ amax, const int* __restrict__ children,
                        const int* __restrict__ parent, const int* __restrict__ leaf_parent,
                        float4* node_min, float4* node_max, int* flags, int ns, int nq, BvhNode* __restrict__ nodes) {
    int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    int node = leaf_parent[pos];
    bool first = true;
    while (node >= 0) {
        if (!PRT_CHECK(node < n - 1, PRT_CHK_REFIT)) return;
        // topology (written by k_karras, read-only here) is fetched BEFORE the fence/atomic so that its latency overlaps
        // theirs: per level the dependent chain is atomic -> child boxes, not atomic -> children -> boxes -> parent
        const int2 ch = __ldg(reinterpret_cast<const int2*>(children) + node);
        const int up = __ldg(&parent[node]);
        if (first && (ch.x & ch.y) < 0) {
            // both children are leaves (only possible on a thread's first level): the left leaf's thread merges, the right
            // one stops - no counter, no fences, the leaf boxes were written by an earlier kernel
            if (pos == ~ch.y) return;
        } else {
            if (!first) __threadfence();                                // publish the box this thread wrote one level below
            if (atomicAdd(&flags[node], 1) == 0) return;
            __threadfence();
        }
        first = false;
        float4 mn[2], mx[2]; int code[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            int c = k ? ch.y : ch.x;
            if (c < 0) { uint32_t q = sorted_prim[~c]; mn[k] = amin[2 * (size_t)~c]; mx[k] = amax[2 * (size_t)~c]; code[k] = leaf_code(~c, q, ns, nq); }
            else { mn[k] = __ldcg(&node_min[2 * c]); mx[k] = __ldcg(&node_max[2 * c]); code[k] = c; }
        }
        BvhNode nd;
        nd.lmin = make_float4(mn[0].x, mn[0].y, mn[0].z, __int_as_float(code[0]));
        nd.lmax = make_float4(mx[0].x, mx[0].y, mx[0].z, __int_as_float(code[1]));
        nd.rmin = make_float4(mn[1].x, mn[1].y, mn[1].z, 0.f);
        nd.rmax = make_float4(mx[1].x, mx[1].y, mx[1].z, 0.f);
        nodes[node] = nd;
        __stcg(&node_min[2 * node], make_float4(fminf(mn[0].x, mn[1].x), fminf(mn[0].y, mn[1].y), fminf(mn[0].z, mn[1].z), 0.f));
        __stcg(&node_max[2 * node], make_float4(fmaxf(mx[0].x, mx[1].x), fmaxf(mx[0].y, mx[1].y), fmaxf(mx[0].z, mx[1].z), 0.f));
        node = up;
    }
}

// triangle planes in leaf-slot order so a leaf code indexes them directly
__global__ void k_gather_tris(int n, int ns, int nq, int tri_prim_base, const uint32_t* __restrict__ sorted_prim,
                              const float4* __restrict__ in_v, const float4* __restrict__ in_n,
                              float4* __restrict__ out_v, float4* __restrict__ out_n,
                              const float4* __restrict__ prim_box, float4* __restrict__ leaf_box, int positions_only) {
    int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    int q = (int)sorted_prim[pos];
    // leaf boxes in SORTED order for the refit (written coalesced here, where all 32 lanes work, instead of gathered there
    // at 5-6 live lanes per warp); a triangle's box comes from the vertices this thread loads anyway - same min/max as k_prim_bounds
    if (q < ns + nq) { leaf_box[2 * (size_t)pos] = prim_box[2 * (size_t)q]; leaf_box[2 * (size_t)pos + 1] = prim_box[2 * (size_t)q + 1]; return; }
    int t = q - ns - nq;
    // one 96-byte input record per triangle (positions, material in v1.w, normals): one random gather instead of three
    float4 a, b, c, n0, n1, n2;                                          // 3 x 256-bit loads of the 96-byte record
    if (positions_only) {                                                // normals / materials follow in k_finish_tris
        ldg8(&in_v[6 * (size_t)t], a, b); c = __ldg(&in_v[6 * (size_t)t + 2]);
        n0 = n1 = n2 = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
        ldg8(&in_v[6 * (size_t)t], a, b); ldg8(&in_v[6 * (size_t)t + 2], c, n0); ldg8(&in_v[6 * (size_t)t + 4], n1, n2);
    }
    leaf_box[2 * (size_t)pos] = make_float4(fminf(fminf(a.x, b.x), c.x), fminf(fminf(a.y, b.y), c.y), fminf(fminf(a.z, b.z), c.z), 0.f);
    leaf_box[2 * (size_t)pos + 1] = make_float4(fmaxf(fmaxf(a.x, b.x), c.x), fmaxf(fmaxf(a.y, b.y), c.y), fmaxf(fmaxf(a.z, b.z), c.z), 0.f);
    a.w = __int_as_float(tri_prim_base + t); c.w = 0.f;
    out_v[3 * pos] = a; out_v[3 * pos + 1] = b; out_v[3 * pos + 2] = c;
    if (positions_only) return;
    out_n[3 * pos] = n0; out_n[3 * pos + 1] = n1; out_n[3 * pos + 2] = n2;
}

// Second half of a positions-only build (prt_set_scene_build): the vertex normals and material indices, which crossed PCIe
// while the hierarchy was being built, go to their leaf slots - the same values k_ingest_tris + k_gather_tris write.
__global__ void k_finish_tris(int n, int ns, int nq, const uint32_t* __restrict__ sorted_prim, const float* __restrict__ nrm,
                              const int* __restrict__ mat, int mat_default, float4* __restrict__ out_v, float4* __restrict__ out_n) {
    int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    int q = (int)sorted_prim[pos];
    if (q < ns + nq) return;
    size_t t = (size_t)(q - ns - nq);
    out_v[3 * pos + 1].w = __int_as_float(mat ? mat[t] : mat_default);
    if (nrm) {
#pragma unroll
        for (int k = 0; k < 3; ++k) out_n[3 * pos + k] = make_float4(nrm[9 * t + 3 * k], nrm[9 * t + 3 * k + 1], nrm[9 * t + 3 * k + 2], 0.f);
    } else {                                                              // flat normal from the gathered vertices: k_ingest_tris' arithmetic
        const float4 a = out_v[3 * pos], b = out_v[3 * pos + 1], c = out_v[3 * pos + 2];
        float e1x = b.x - a.x, e1y = b.y - a.y, e1z = b.z - a.z, e2x = c.x - a.x, e2y = c.y - a.y, e2z = c.z - a.z;
        float nx = e1y * e2z - e2y * e1z, ny = e1z * e2x - e2z * e1x, nz = e1x * e2y - e2x * e1y;
#pragma unroll
        for (int k = 0; k < 3; ++k) out_n[3 * pos + k] = make_float4(nx, ny, nz, 0.f);
    }
}
int lbvh_finish_tris(Ctx* c) {
    const int n = c->n_prims;
    if (n <= 0 || c->n_tris <= 0 || !c->tri_deferred) return PRT_OK;
    k_finish_tris<<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->n_spheres, c->n_quads, (const uint32_t*)c->vals[c->sorted_slot].p,
        c->def_normals ? (const float*)c->raw_n.p : nullptr, c->def_mat ? (const int*)c->raw_m.p : nullptr, c->def_mat_default,
        (float4*)c->tri_v.p, (float4*)c->tri_n.p);
    return PRT_OK;
}

// ------------------------------------------------------------------ PLOC: the tree-quality pass
// Parallel locally-ordered clustering (Meister & Bittner 2018) over the Morton-sorted leaves - SURVEY 8(f)-4, the BVH
// quality step the reference's own notes ask for (wavefront.md:86-96). The Karras hierarchy splits by code bits only: one
// large primitive (a ground quad) inflates every ancestor on its root path. PLOC builds bottom-up by box area instead:
// every round, each cluster looks `R` positions left and right in the current array, picks the neighbour whose union box
// has the smallest half-area (ties: position i^1 if it is among them - runs of identical boxes then pair up instead of
// forming a chain -, else the lowest position; the globally closest pair is always mutual); mutual nearest neighbours merge; the array is compacted, order kept.
// Measured on the CPU restatement (oracle/prt_oracle.cpp:build_ploc, same bits): wide-node visits per bounce ray
// dragon 8.0 -> 4.9, 809 spheres 9.4 -> 7.0, uniform triangle soup unchanged (79.2 -> 79.9).
//
// ONE persistent cooperative kernel runs every round: tiles of 1024 clusters are handed out by ticket; a tile loads its
// clusters plus a halo of 2R into shared memory, finds nearest neighbours, and compacts through a chained scan with
// decoupled look-back over (clusters kept, merges) - so output positions and node numbers need no second pass. Node
// ids count DOWN from n-2 in merge order, the last merge (the root) is node 0 like Karras's numbering; the merge writes
// the finished 64-byte traversal node (both child boxes) - there is no refit pass. One grid.sync() per round; once a
// round fits one tile, CTA 0 finishes alone with __syncthreads(). Everything another CTA wrote is read with ld.cg.
constexpr int PL_THREADS = 256, PL_ITEMS = 4, PL_TILE = PL_THREADS * PL_ITEMS, PL_RMAX = 8;
constexpr int PL_SPAN = PL_TILE + 4 * PL_RMAX;
constexpr int PL_LOAD_TRIPS = (PL_SPAN + PL_THREADS - 1) / PL_THREADS;
constexpr int PL_SMALL_ROUND = 160 << 10;   // rounds with at most this many clusters use 256-cluster tiles (one per thread)
constexpr int PL_MAX_ROUNDS = 192;          // a round may merge a single pair (nested boxes): give up, Karras takes over
constexpr int PL_MAX_HEIGHT = 84;           // 3 * height + 1 <= 256 stack entries (wavefront.cu STACK_TOTAL)
// ctl words
enum { PC_COUNT = 0, PC_NEXT = 2, PC_TICKET = 4, PC_HEIGHT = 6, PC_ROUNDS = 7, PC_FAIL = 8, PC_BAR_COUNT = 9, PC_BAR_GEN = 10, PC_WORDS = 16,
       PC_TRACE = 16, PC_ALL = PC_TRACE + 2 * (PL_MAX_ROUNDS + 1) };   // per round: cluster count, %globaltimer (ns, low word) at its start

// Tile status words of the chained scan carry their own payload (flag and counts in ONE 64-bit word), so relaxed
// device-scope accesses suffice: nothing else is read on the strength of having seen a flag.
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v; asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// half-area of the union of two boxes; the exact expression is part of the bit-exact contract with the oracle
__device__ __forceinline__ float union_area(float4 a0, float4 a1, float4 b0, float4 b1) {
    const float dx = fmaxf(a1.x, b1.x) - fminf(a0.x, b0.x);
    const float dy = fmaxf(a1.y, b1.y) - fminf(a0.y, b0.y);
    const float dz = fmaxf(a1.z, b1.z) - fminf(a0.z, b0.z);
    return (dx * dy + dy * dz) + dz * dx;
}

// RT = search radius known at compile time (4, the default: the neighbour loop is unrolled) or 0 = the run-time value R.
template <int RT>
__global__ void __launch_bounds__(PL_THREADS)
k_ploc(int n, int R_, int ns, int nq, const uint32_t* __restrict__ sorted_prim, const float4* leaf_box,
       int* code_a, int* code_b, float4* box_a, float4* box_b, unsigned long long* status, int* ctl,
       BvhNode* __restrict__ nodes, int* __restrict__ children, int* __restrict__ parent) {
    cg::grid_group grid = cg::this_grid();
    __shared__ float4 s_min[PL_SPAN], s_max[PL_SPAN];
    __shared__ int s_code[PL_SPAN], s_nn[PL_SPAN];
    __shared__ uint32_t s_warp[PL_THREADS / 32];
    __shared__ int s_tile;
    __shared__ uint32_t s_excl[2];
    const int R = RT ? RT : R_;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const float INF = __int_as_float(0x7f800000);
    for (int round = 0;; ++round) {
        const int par = round & 1;
        const int c = __ldcg(&ctl[PC_COUNT + par]), next = __ldcg(&ctl[PC_NEXT + par]);
        if (blockIdx.x == 0 && tid == 0) {                   // round trace (prt_debug_ploc_trace): one store pair per round
            unsigned long long ns_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ns_));
            ctl[PC_TRACE + 2 * round] = c; ctl[PC_TRACE + 2 * round + 1] = (int)(uint32_t)ns_;
        }
        if (c <= 1) break;
        if (round >= PL_MAX_ROUNDS) { if (blockIdx.x == 0 && tid == 0) ctl[PC_FAIL] = 1; break; }
        const bool single = c <= PL_TILE;                    // the rest of the build fits one tile: CTA 0 finishes alone
        if (single && blockIdx.x != 0) break;
        // clusters per thread: 4 while there is work for every CTA, 1 for the thin middle rounds (a round is then as long as
        // ONE tile's critical path - load, search, scan, look-back, write -, so short tiles on many CTAs beat long ones on few)
        const int ipt = (single || c > PL_SMALL_ROUND) ? PL_ITEMS : 1;
        const int tile = PL_THREADS * ipt;
        const int ntiles = (c + tile - 1) / tile;
        // round r reads what round r-1 wrote (round 0: the sorted leaf boxes, codes implicit) and writes the other buffer
        const float4* bin = round == 0 ? leaf_box : (par ? box_b : box_a);
        const int* cin = par ? code_b : code_a;
        float4* bout = par ? box_a : box_b;
        int* cout = par ? code_a : code_b;
        const unsigned long long tag = (unsigned long long)(round & 63) << 58;
        for (;;) {
            __syncthreads();
            if (tid == 0) s_tile = single ? 0 : atomicAdd(&ctl[PC_TICKET + par], 1);
            __syncthreads();
            const int t = s_tile;
            if (t >= ntiles) break;
            const int t0 = t * tile, lo = t0 - 2 * R;        // global position of shared slot 0
            // The tile's boxes go global -> shared with cp.async (L2 only: other CTAs wrote them last round), all in flight at
            // once; as a load/store loop each of its five trips waited for its own DRAM round trip before the next began.
            {
                int codes[PL_LOAD_TRIPS];
#pragma unroll
                for (int it = 0; it < PL_LOAD_TRIPS; ++it) {
                    const int k = tid + it * PL_THREADS, g = lo + k;
                    codes[it] = 0;
                    if (k < tile + 4 * R && g >= 0 && g < c) {
                        const unsigned dmin = (unsigned)__cvta_generic_to_shared(&s_min[k]), dmax = (unsigned)__cvta_generic_to_shared(&s_max[k]);
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dmin), "l"(bin + 2 * (size_t)g) : "memory");
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dmax), "l"(bin + 2 * (size_t)g + 1) : "memory");
                        codes[it] = round == 0 ? ~g : __ldcg(&cin[g]);
                    }
                }
#pragma unroll
                for (int it = 0; it < PL_LOAD_TRIPS; ++it) {
                    const int k = tid + it * PL_THREADS;
                    if (k < tile + 4 * R) s_code[k] = codes[it];
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
                asm volatile("cp.async.wait_group 0;" ::: "memory");
            }
            __syncthreads();
            // nearest neighbour of every cluster of the tile and of R more on either side (their choice decides "mutual")
            for (int k = R + tid; k < tile + 3 * R; k += PL_THREADS) {
                const int g = lo + k;
                int bj = -1;
                if (g >= 0 && g < c) {
                    const float4 a0 = s_min[k], a1 = s_max[k];
                    float best = INF;
                    if (RT) {                                // ascending j, like the run-time loop: ties go to the lowest position
#pragma unroll
                        for (int dr = -RT; dr <= RT; ++dr) {
                            if (dr == 0) continue;
                            const int j = g + dr;
                            const float ar = (j >= 0 && j < c) ? union_area(a0, a1, s_min[k + dr], s_max[k + dr]) : INF;
                            if ((ar < best || (ar == best && j == (g ^ 1))) && j >= 0 && j < c) { best = ar; bj = j; }
                        }
                    } else {
                        const int j0 = max(0, g - R), j1 = min(c - 1, g + R);
                        for (int j = j0; j <= j1; ++j) {
                            if (j == g) continue;
                            const float ar = union_area(a0, a1, s_min[j - lo], s_max[j - lo]);
                            if (ar < best || (ar == best && j == (g ^ 1))) { best = ar; bj = j; }   // ties: the partner g^1, else the lowest position
                        }
                    }
                }
                s_nn[k] = bj;
            }
            __syncthreads();
            // `ipt` consecutive clusters per thread: kept (not absorbed by a lower position) / merging (lower one of a mutual pair)
            uint32_t flags = 0, cnt = 0;                     // cnt: kept in the low half, merges in the high half
#pragma unroll
            for (int m = 0; m < PL_ITEMS; ++m) {
                const int k = 2 * R + tid * ipt + m, g = lo + k;
                if (m < ipt && g < c) {
                    const int j = s_nn[k];
                    const bool mutual = s_nn[j - lo] == g;
                    const bool keep = !mutual || g < j, merge = mutual && g < j;
                    flags |= (keep ? 1u : 0u) << (2 * m) | (merge ? 2u : 0u) << (2 * m);
                    cnt += (keep ? 1u : 0u) + (merge ? 0x10000u : 0u);
                }
            }
            uint32_t inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
            if (lane == 31) s_warp[wid] = inc;
            __syncthreads();
            uint32_t wbase = 0, total = 0;
#pragma unroll
            for (int k = 0; k < PL_THREADS / 32; ++k) { const uint32_t v = s_warp[k]; wbase += k < wid ? v : 0u; total += v; }
            uint32_t run = wbase + inc - cnt;                // exclusive prefix inside the tile
            // Chained scan across tiles (tickets are taken in order, so every predecessor is resident and will publish): warp 0
            // publishes this tile's aggregate, then looks back 128 tiles per step - four independent loads per lane in flight -
            // until it meets an inclusive prefix. (With ~600 tiles in flight everybody waits on everybody, so the walk IS the
            // tile's critical path: one thread, one tile per step took the whole build; 32 per step still cost 13 us per tile.)
            if (wid == 0) {
                const unsigned long long agg = ((unsigned long long)(total >> 16) << 28) | (unsigned long long)(total & 0xffffu);
                const unsigned long long PAY = (1ull << 56) - 1ull;
                unsigned long long excl = 0;
                if (t > 0) {
                    if (lane == 0) st_relaxed_u64(&status[t], tag | (1ull << 56) | agg);
                    bool done = false;
                    for (int base = t - 1; !done; base -= 128) {
                        unsigned long long v[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) { const int p = base - 32 * u - lane; v[u] = p >= 0 ? ld_relaxed_u64(&status[p]) : 0ull; }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            if (done) break;
                            const int p = base - 32 * u - lane;
                            if (p >= 0) while ((v[u] >> 58) != (tag >> 58) || ((v[u] >> 56) & 3ull) == 0ull) v[u] = ld_relaxed_u64(&status[p]);
                            const bool stop = p < 0 || ((v[u] >> 56) & 3ull) == 2ull;   // an inclusive prefix, or nothing before tile 0
                            const uint32_t ball = __ballot_sync(0xffffffffu, stop);
                            const int first = ball ? __ffs(ball) - 1 : 32;
                            unsigned long long part = (p >= 0 && lane <= first) ? (v[u] & PAY) : 0ull;
#pragma unroll
                            for (int o = 16; o > 0; o >>= 1) part += __shfl_down_sync(0xffffffffu, part, o);
                            excl += __shfl_sync(0xffffffffu, part, 0);
                            done = ball != 0u;
                        }
                    }
                }
                if (lane == 0) {
                    st_relaxed_u64(&status[t], tag | (2ull << 56) | (excl + agg));
                    s_excl[0] = (uint32_t)(excl & ((1ull << 28) - 1ull)); s_excl[1] = (uint32_t)(excl >> 28);
                    if (t == ntiles - 1) {                   // the last tile knows the totals: parameters of the next round
                        const unsigned long long all = excl + agg;
                        const int c_next = (int)(all & ((1ull << 28) - 1ull)), merges = (int)(all >> 28);
                        ctl[PC_COUNT + (par ^ 1)] = c_next; ctl[PC_NEXT + (par ^ 1)] = next - merges;
                        ctl[PC_TICKET + (par ^ 1)] = 0; ctl[PC_ROUNDS] = round + 1;
                    }
                }
            }
            __syncthreads();
            const uint32_t E = s_excl[0], M = s_excl[1];
#pragma unroll
            for (int m = 0; m < PL_ITEMS; ++m) {
                const uint32_t f = (flags >> (2 * m)) & 3u;
                if (!(f & 1u)) continue;
                const int k = 2 * R + tid * ipt + m;
                const size_t pos = (size_t)E + (run & 0xffffu);
                if (!PRT_CHECK(pos < (size_t)c && k < PL_SPAN, PRT_CHK_PLOC)) continue;
                if (f & 2u) {
                    const int id = next - 1 - (int)(M + (run >> 16));
                    const int kj = s_nn[k] - lo;
                    if (!PRT_CHECK(id >= 0 && id < n - 1 && kj >= 0 && kj < PL_SPAN, PRT_CHK_PLOC)) continue;
                    const float4 a0 = s_min[k], a1 = s_max[k], b0 = s_min[kj], b1 = s_max[kj];
                    const int ci = s_code[k], cj = s_code[kj];
                    const int h = max(__float_as_int(a0.w), __float_as_int(b0.w)) + 1;
                    const int qi = (ci < 0 && ns + nq) ? (int)sorted_prim[~ci] : ns + nq, qj = (cj < 0 && ns + nq) ? (int)sorted_prim[~cj] : ns + nq;
                    BvhNode nd;
                    nd.lmin = make_float4(a0.x, a0.y, a0.z, __int_as_float(ci < 0 ? leaf_code(~ci, (uint32_t)qi, ns, nq) : ci));
                    nd.lmax = make_float4(a1.x, a1.y, a1.z, __int_as_float(cj < 0 ? leaf_code(~cj, (uint32_t)qj, ns, nq) : cj));
                    nd.rmin = make_float4(b0.x, b0.y, b0.z, 0.f);
                    nd.rmax = make_float4(b1.x, b1.y, b1.z, 0.f);
                    stg8(&nodes[id].lmin, nd.lmin, nd.lmax); stg8(&nodes[id].rmin, nd.rmin, nd.rmax);
                    *reinterpret_cast<int2*>(children + 2 * (size_t)id) = make_int2(ci, cj);
                    if (ci >= 0) parent[ci] = id;
                    if (cj >= 0) parent[cj] = id;
                    if (id == 0) { parent[0] = -1; ctl[PC_HEIGHT] = h; }
                    __stcg(&cout[pos], id);
                    stg8(&bout[2 * pos], make_float4(fminf(a0.x, b0.x), fminf(a0.y, b0.y), fminf(a0.z, b0.z), __int_as_float(h)),
                         make_float4(fmaxf(a1.x, b1.x), fmaxf(a1.y, b1.y), fmaxf(a1.z, b1.z), 0.f));
                    run += 0x10000u;
                } else {
                    __stcg(&cout[pos], s_code[k]); stg8(&bout[2 * pos], s_min[k], s_max[k]);
                }
                run += 1u;
            }
            if (single) break;
        }
        // (a hand-rolled counter/generation barrier was slower than cg's: 18 vs 10 us per thin round with 1184 CTAs of 128 threads)
        if (single) { __threadfence(); __syncthreads(); } else grid.sync();
    }
}

// ------------------------------------------------------------------ 4-wide quantised nodes
// Collapse the two-child tree greedily by SURFACE AREA: a wide node starts with its two children and, twice, replaces the
// internal slot with the largest box by that slot's two children (the new pair takes the slot's place, order kept), so the
// 2..4 slots may sit one to three levels below. Unbalanced LBVH subtrees (a big primitive next to small ones) get their
// big side opened instead of a fixed two levels. EVERY internal node gets a record, at its own index - which ones a ray
// ever visits follows from the child codes, the root is still node 0, child codes need no remapping.
// grid step of a wide node on one axis: the smallest power of two s with ext / s <= 253 (one step below the boxes stays
// free, see k_emit_wide), at least 2 ulp of the larger bound (so that origin = mn - s is a different float), >= 2^-60
__device__ __forceinline__ float pow2_ceil_scale(float ext, float mn, float mx) {
    int e = -60;
    if (ext > 0.0f) { int e1; frexpf(ext * (1.0f / 253.0f), &e1); e = e1 > e ? e1 : e; }   // ext/253 = m * 2^e1, m in [0.5, 1)
    const float m = fmaxf(fabsf(mn), fabsf(mx));
    if (m > 0.0f) { int e2; frexpf(m, &e2); e2 -= 23; e = e2 > e ? e2 : e; }
    return ldexpf(1.0f, e);
}

__device__ __forceinline__ float box_area(const float* lo, const float* hi) {
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return (dx * dy + dy * dz) + dz * dx;
}

// slot S of a wide node under construction is replaced by the two children `ch` of the node it held; the slots behind
// it move up by one. S is a template parameter so that every array index is a compile-time constant (registers).
template <int S>
__device__ __forceinline__ void open_slot(float (&lo)[4][3], float (&hi)[4][3], int (&code)[4], int ns, const BvhNode& ch) {
#pragma unroll
    for (int c = 3; c > S + 1; --c)
        if (c <= ns) {
            code[c] = code[c - 1];
#pragma unroll
            for (int k = 0; k < 3; ++k) { lo[c][k] = lo[c - 1][k]; hi[c][k] = hi[c - 1][k]; }
        }
    lo[S][0] = ch.lmin.x; lo[S][1] = ch.lmin.y; lo[S][2] = ch.lmin.z; hi[S][0] = ch.lmax.x; hi[S][1] = ch.lmax.y; hi[S][2] = ch.lmax.z;
    lo[S + 1][0] = ch.rmin.x; lo[S + 1][1] = ch.rmin.y; lo[S + 1][2] = ch.rmin.z;
    hi[S + 1][0] = ch.rmax.x; hi[S + 1][1] = ch.rmax.y; hi[S + 1][2] = ch.rmax.z;
    code[S] = __float_as_int(ch.lmin.w); code[S + 1] = __float_as_int(ch.lmax.w);
}

__global__ void k_emit_wide(int n_internal, const int* __restrict__ parent, const BvhNode* __restrict__ nodes,
                            WideNode* __restrict__ wide) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_internal) return;
    (void)parent;
    float lo[4][3], hi[4][3]; int code[4]; int ns = 2;
    {
        BvhNode nd;                                                      // 2 x 256-bit loads: every 32-byte sector is requested once
        ldg8(&nodes[i].lmin, nd.lmin, nd.lmax); ldg8(&nodes[i].rmin, nd.rmin, nd.rmax);
        lo[0][0] = nd.lmin.x; lo[0][1] = nd.lmin.y; lo[0][2] = nd.lmin.z; hi[0][0] = nd.lmax.x; hi[0][1] = nd.lmax.y; hi[0][2] = nd.lmax.z;
        lo[1][0] = nd.rmin.x; lo[1][1] = nd.rmin.y; lo[1][2] = nd.rmin.z; hi[1][0] = nd.rmax.x; hi[1][1] = nd.rmax.y; hi[1][2] = nd.rmax.z;
        code[0] = __float_as_int(nd.lmin.w); code[1] = __float_as_int(nd.lmax.w);
    }
    // every slot index below is a compile-time constant (open_slot<S>), so the four boxes stay in registers - indexed by
    // the chosen slot they lived in a 112-byte local-memory frame (12 GB of L2-level traffic on 10 M triangles)
#pragma unroll
    for (int round = 0; round < 2; ++round) {
        int s = -1; float best = -1.0f;
#pragma unroll
        for (int c = 0; c < 3; ++c)
            if (c < ns && code[c] >= 0) { const float ar = box_area(lo[c], hi[c]); if (ar > best) { best = ar; s = c; } }   // ties: lowest slot
        if (s >= 0) {
            const int cs = s == 0 ? code[0] : s == 1 ? code[1] : code[2];
            if (!PRT_CHECK(cs < n_internal, PRT_CHK_REFIT)) break;
            BvhNode ch;
            ldg8(&nodes[cs].lmin, ch.lmin, ch.lmax); ldg8(&nodes[cs].rmin, ch.rmin, ch.rmax);
            if (s == 0) open_slot<0>(lo, hi, code, ns, ch);
            else if (s == 1) open_slot<1>(lo, hi, code, ns, ch);
            else open_slot<2>(lo, hi, code, ns, ch);
            ++ns;
        }
    }
    WideNode w;
    float org[3], sc[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float mn = lo[0][k], mx = hi[0][k];
#pragma unroll
        for (int c = 1; c < 4; ++c) if (c < ns) { mn = fminf(mn, lo[c][k]); mx = fmaxf(mx, hi[c][k]); }
        sc[k] = pow2_ceil_scale(mx - mn, mn, mx);
        // the grid starts ONE STEP BELOW the lowest box: a box that starts at the node's own minimum floors to step 0 with
        // >= half a step of margin instead of being clamped there with none (the fused decode of the walk rounds by up to
        // 2^-9 step: without the margin, rays grazing that corner missed the child - advisor finding, round 1)
        org[k] = mn - sc[k];
        const float rs = 1.0f / sc[k];                                  // sc is a power of two in [2^-60, 2^121]: x * rs == x / sc bit for bit
        uint32_t ql = 0, qh = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t a = 255u, b = 0u;                                  // empty slot: inverted box
            if (c < ns) {
                float fl = floorf((lo[c][k] - org[k]) * rs - 0.015625f), fh = ceilf((hi[c][k] - org[k]) * rs + 0.015625f);
                a = (uint32_t)fminf(fmaxf(fl, 0.0f), 255.0f); b = (uint32_t)fminf(fmaxf(fh, 0.0f), 255.0f);
            }
            ql |= a << (8 * c); qh |= b << (8 * c);
        }
        w.qlo[k] = ql; w.qhi[k] = qh;
    }
    w.ox = org[0]; w.oy = org[1]; w.oz = org[2];
    w.sx = sc[0] * 32768.0f; w.sy = sc[1] * 32768.0f; w.sz = sc[2] * 32768.0f;
#pragma unroll
    for (int c = 0; c < 4; ++c) w.code[c] = c < ns ? code[c] : code[0];
    stg8(&wide[i], make_float4(w.ox, w.oy, w.oz, w.sx),
         make_float4(__int_as_float(w.code[0]), __int_as_float(w.code[1]), __int_as_float(w.code[2]), __int_as_float(w.code[3])));
    stg8(reinterpret_cast<char*>(&wide[i]) + 32, make_float4(__uint_as_float(w.qlo[0]), __uint_as_float(w.qlo[1]), __uint_as_float(w.qlo[2]), __uint_as_float(w.qhi[0])),
         make_float4(__uint_as_float(w.qhi[1]), __uint_as_float(w.qhi[2]), w.sy, w.sz));
}

unsigned lbvh_check_flags() {
#if PRT_CHECKED
    unsigned v = 0, z = 0;
    cudaMemcpyFromSymbol(&v, g_prt_check_flags, sizeof v);
    cudaMemcpyToSymbol(g_prt_check_flags, &z, sizeof z);
    return v;
#else
    return 0u;
#endif
}

// ------------------------------------------------------------------ host driver
// ------------------------------------------------------------------ compact wide nodes (option wide_compact)
// k_emit_wide writes a record for EVERY internal node (any of them could be a wide root); the walk only ever reaches the ones a
// reachable record names as a slot - about 45 % - so half of every 128-byte line of that array is dead weight in L1 and L2.
// One breadth-first pass from the root copies the reachable records into a dense array: level after level, in frontier order,
// a node's internal children get consecutive new indices (siblings share cache lines) and the parent's copy is written with
// the new indices in place. Deterministic: every CTA owns a contiguous piece of the frontier, counts its children (pass A),
// and after a grid.sync takes the sum of the CTAs before it as its base (pass B) - no atomics, the same layout on every run
// and every device. Cooperative launch, two grid.sync per level; frontier entries are (old index, new index) pairs.
constexpr int CW_T = 256;
__device__ __forceinline__ int cw_kids(const WideNode& w, int kid[4]) {
    int nk = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const bool empty = ((w.qlo[0] >> (8 * k)) & 255u) == 255u && ((w.qhi[0] >> (8 * k)) & 255u) == 0u;
        kid[k] = (!empty && w.code[k] >= 0) ? 1 : 0;
        nk += kid[k];
    }
    return nk;
}
__device__ __forceinline__ int cw_block_sum(int v, int* s_w) {          // all CW_T threads; returns the sum to every thread
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    __syncthreads();
    if ((threadIdx.x & 31u) == 0u) s_w[threadIdx.x >> 5] = v;
    __syncthreads();
    int t = 0;
#pragma unroll
    for (int k = 0; k < CW_T / 32; ++k) t += s_w[k];
    return t;
}
__global__ void __launch_bounds__(CW_T) k_compact_wide(const WideNode* __restrict__ wide, WideNode* __restrict__ out, int* __restrict__ qa, int* __restrict__ qb,
                                                       int* __restrict__ cta_tot /* 2 x gridDim: children per CTA, double-buffered by level */) {
    cg::grid_group grid = cg::this_grid();
    __shared__ int s_w[CW_T / 32];
    const int B = (int)gridDim.x, b = (int)blockIdx.x, t = (int)threadIdx.x, lane = t & 31, wid = t >> 5;
    if (b == 0 && t == 0) { qa[0] = 0; qa[1] = 0; }
    grid.sync();
    int cnt = 1, placed = 1;                                             // the same on every CTA: all of them add up the same totals
    int* fa = qa; int* fb = qb;
    for (int level = 0; cnt > 0; ++level) {
        int* tot = cta_tot + (level & 1) * B;
        const int chunk = (((cnt + B - 1) / B + CW_T - 1) / CW_T) * CW_T;
        const int lo = min(cnt, b * chunk), hi = min(cnt, lo + chunk);
        int my = 0;                                                      // pass A: internal children of my piece of the frontier
        for (int i = lo + t; i < hi; i += CW_T) { int kid[4]; const WideNode w = wide[fa[2 * i]]; my += cw_kids(w, kid); }
        my = cw_block_sum(my, s_w);
        if (t == 0) tot[b] = my;
        grid.sync();
        int before = 0, all = 0;                                         // pass B: my base = the children of the CTAs before me
        for (int k = t; k < B; k += CW_T) { const int v = __ldcg(&tot[k]); all += v; if (k < b) before += v; }
        before = cw_block_sum(before, s_w);
        all = cw_block_sum(all, s_w);
        int run = before;
        for (int i0 = lo; i0 < hi; i0 += CW_T) {
            const int i = i0 + t;
            const bool act = i < hi;
            WideNode w; int kid[4] = {0, 0, 0, 0}; int nk = 0, new_i = 0;
            if (act) { w = wide[fa[2 * i]]; new_i = fa[2 * i + 1]; nk = cw_kids(w, kid); }
            int incl = nk;                                               // block-wide exclusive prefix of nk
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
            __syncthreads();
            if (lane == 31) s_w[wid] = incl;
            __syncthreads();
            int wbase = 0, ctotal = 0;
#pragma unroll
            for (int k = 0; k < CW_T / 32; ++k) { const int v = s_w[k]; if (k < wid) wbase += v; ctotal += v; }
            if (act) {
                int pos = run + wbase + incl - nk;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (kid[k]) { fb[2 * pos] = w.code[k]; fb[2 * pos + 1] = placed + pos; w.code[k] = placed + pos; ++pos; }
#pragma unroll
                for (int k = 1; k < 4; ++k) {                            // empty slots keep "the code of slot 0" (see WideNode)
                    const bool empty = ((w.qlo[0] >> (8 * k)) & 255u) == 255u && ((w.qhi[0] >> (8 * k)) & 255u) == 0u;
                    if (empty) w.code[k] = w.code[0];
                }
                out[new_i] = w;
            }
            run += ctotal;
        }
        placed += all; cnt = all;
        int* x = fa; fa = fb; fb = x;
        grid.sync();
    }
    if (b == 0 && t == 0) cta_tot[2 * B] = placed;                       // records in the dense array
}

int lbvh_build(Ctx* c) {
    const int n = c->n_prims, ns = c->n_spheres, nq = c->n_quads, nt = c->n_tris;
    c->bvh_built = false; c->wide_built = false;
    cudaStream_t st = c->stream;
    int rc;
    if ((rc = dev_reserve(c, c->scene_bounds, 12 * sizeof(int)))) return rc;
    if (n == 0) {
        k_init_bounds<<<1, 32, 0, st>>>((int*)c->scene_bounds.p);
        c->root_code = 0; c->bvh_built = true; c->build_ms = 0.f;
        PRT_CUDA(c, cudaStreamSynchronize(st));
        return PRT_OK;
    }
    const size_t nn = (size_t)n;
    if ((rc = dev_reserve(c, c->aabb_min, nn * 32)) || (rc = dev_reserve(c, c->leaf_box, nn * 32)) ||
        (rc = dev_reserve(c, c->morton, nn * 4)) ||
        (rc = dev_reserve(c, c->keys[0], nn * 4)) || (rc = dev_reserve(c, c->keys[1], nn * 4)) ||
        (rc = dev_reserve(c, c->vals[0], nn * 4)) || (rc = dev_reserve(c, c->vals[1], nn * 4)) ||
        (rc = dev_reserve(c, c->children, nn * 8)) || (rc = dev_reserve(c, c->parent, nn * 8)) ||
        (rc = dev_reserve(c, c->node_min, nn * 32)) ||
        (rc = dev_reserve(c, c->flags, nn * 4)) || (rc = dev_reserve(c, c->nodes, nn * sizeof(BvhNode))) ||
        (rc = dev_reserve(c, c->tri_v, nn * 48)) || (rc = dev_reserve(c, c->tri_n, nn * 48)))
        return rc;
    const int nblocks = (n + RS_TILE - 1) / RS_TILE;
    if ((rc = dev_reserve(c, c->block_hist, ((size_t)nblocks * 256 + 256) * 4))) return rc;   // + 256 digit totals
    const bool want_wide = n > 1 && (c->bvh_wide > 0 || (c->bvh_wide < 0 && n >= c->wide_min_prims));
    if (want_wide && (rc = dev_reserve(c, c->wnodes, nn * sizeof(WideNode)))) return rc;
    const bool want_ploc = n > 2 && c->bvh_builder != 0;
    if (want_ploc && ((rc = dev_reserve(c, c->ploc_code, nn * 8)) || (rc = dev_reserve(c, c->ploc_box, nn * 32)) ||
                      (rc = dev_reserve(c, c->ploc_status, std::max<size_t>((nn + PL_TILE - 1) / PL_TILE, PL_SMALL_ROUND / PL_THREADS + 1) * 8)) ||
                      (rc = dev_reserve(c, c->ploc_ctl, PC_ALL * 4)))) return rc;
    // every allocation is done: the timed region below is kernels only (cudaMalloc synchronises and is not build work)
    PRT_CUDA(c, cudaEventRecord(c->ev0, st));
    k_init_bounds<<<1, 32, 0, st>>>((int*)c->scene_bounds.p);

    const int T = 256, G = (n + T - 1) / T;
    // boxes of the spheres and quads; the triangles' boxes and bounds come from the ingest pass of prt_set_scene
    k_prim_bounds<<<std::max(1, (ns + nq + T - 1) / T), T, 0, st>>>((const SphereRec*)c->spheres.p, ns, (const QuadRec*)c->quads.p, nq,
                                   (const float4*)c->tri_in_v.p, 0, nt ? (const int*)c->tri_bounds.p : nullptr, (float4*)c->aabb_min.p,
                                   (float4*)c->aabb_min.p + 1, (int*)c->scene_bounds.p);
    k_morton<<<G, T, 0, st>>>((const float4*)c->aabb_min.p, (const float4*)c->aabb_min.p + 1, n,
                              (const int*)c->scene_bounds.p + (c->morton_centroid ? 6 : 0), (uint32_t*)c->keys[0].p,
                              (uint32_t*)c->vals[0].p, (uint32_t*)c->morton.p);
    int cur = 0;
    for (int pass = 0; pass < 4; ++pass) {
        int shift = pass * 8;
        k_rs_hist<<<nblocks, RS_THREADS, 0, st>>>((const uint32_t*)c->keys[cur].p, n, shift,
                                                  (uint32_t*)c->block_hist.p, nblocks);
        k_rs_scan_rows<<<256, 256, 0, st>>>((uint32_t*)c->block_hist.p, nblocks, (uint32_t*)c->block_hist.p + (size_t)nblocks * 256);
        k_rs_scatter<<<nblocks, RS_THREADS, 0, st>>>((const uint32_t*)c->keys[cur].p, (const uint32_t*)c->vals[cur].p,
                                                     (uint32_t*)c->keys[cur ^ 1].p, (uint32_t*)c->vals[cur ^ 1].p,
                                                     n, shift, (const uint32_t*)c->block_hist.p, nblocks,
                                                     (const uint32_t*)c->block_hist.p + (size_t)nblocks * 256);
        cur ^= 1;
    }
    c->sorted_slot = cur;
    const uint32_t* skeys = (const uint32_t*)c->keys[cur].p;
    const uint32_t* sprim = (const uint32_t*)c->vals[cur].p;
    int* parent = (int*)c->parent.p;
    int* leaf_parent = parent + n;
    c->builder_used = 0; c->bvh_height = 0; c->ploc_rounds = 0;
    if (n > 1) {
        k_gather_tris<<<G, T, 0, st>>>(n, ns, nq, c->tri_prim_base, sprim, (const float4*)c->tri_in_v.p,
                                       (const float4*)c->tri_in_v.p + 3, (float4*)c->tri_v.p, (float4*)c->tri_n.p,
                                       (const float4*)c->aabb_min.p, (float4*)c->leaf_box.p, c->tri_deferred ? 1 : 0);
        bool ploc_done = false;
        if (want_ploc) {
            // PLOC over the sorted leaves; on failure (no convergence / a tree deeper than the traversal stack allows) the
            // Karras hierarchy below is built instead - it reads only the sorted keys and leaf boxes, which PLOC leaves alone
            int* ctl = (int*)c->ploc_ctl.p;
            const int ctl0[PC_WORDS] = {n, 0, n - 1, 0};
            PRT_CUDA(c, cudaMemcpyAsync(ctl, ctl0, sizeof ctl0, cudaMemcpyHostToDevice, st));
            const int ntiles = (n + PL_TILE - 1) / PL_TILE;
            PRT_CUDA(c, cudaMemsetAsync(c->ploc_status.p, 0, c->ploc_status.bytes, st));
            int per_sm = 0;
            int R = std::min(PL_RMAX, std::max(1, c->ploc_radius));
            const void* kern = R == 4 ? (const void*)k_ploc<4> : (const void*)k_ploc<0>;
            PRT_CUDA(c, R == 4 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ploc<4>, PL_THREADS, 0)
                               : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ploc<0>, PL_THREADS, 0));
            const int grid = std::max(1, std::min(ntiles, c->num_sms * std::max(1, per_sm)));
            int n_ = n, ns_ = ns, nq_ = nq;
            const uint32_t* sp_ = sprim; const float4* lb_ = (const float4*)c->leaf_box.p;
            int *ca = (int*)c->ploc_code.p, *cb = ca + n;
            float4 *ba = (float4*)c->ploc_box.p, *bb = (float4*)c->node_min.p;
            unsigned long long* stt = (unsigned long long*)c->ploc_status.p;
            BvhNode* nodes_ = (BvhNode*)c->nodes.p; int* ch_ = (int*)c->children.p;
            void* args[] = {&n_, &R, &ns_, &nq_, &sp_, &lb_, &ca, &cb, &ba, &bb, &stt, &ctl, &nodes_, &ch_, &parent};
            PRT_CUDA(c, cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(PL_THREADS), args, 0, st));
            int* res = c->ploc_trace;
            PRT_CUDA(c, cudaMemcpyAsync(res, ctl, PC_ALL * 4, cudaMemcpyDeviceToHost, st));
            PRT_CUDA(c, cudaStreamSynchronize(st));
            c->ploc_rounds = res[PC_ROUNDS];
            ploc_done = !res[PC_FAIL] && res[PC_HEIGHT] <= PL_MAX_HEIGHT;
            if (ploc_done) { c->builder_used = 1; c->bvh_height = res[PC_HEIGHT]; }
        }
        if (!ploc_done) {
            PRT_CUDA(c, cudaMemsetAsync(c->flags.p, 0, nn * 4, st));
            k_karras<<<G, T, 0, st>>>(skeys, n, (int*)c->children.p, parent, leaf_parent);
            k_refit<<<G, T, 0, st>>>(n, sprim, (const float4*)c->leaf_box.p, (const float4*)c->leaf_box.p + 1,
                                     (const int*)c->children.p, parent, leaf_parent,
                                     (float4*)c->node_min.p, (float4*)c->node_min.p + 1, (int*)c->flags.p, ns, nq, (BvhNode*)c->nodes.p);
        }
        if (want_wide) {
            k_emit_wide<<<G, T, 0, st>>>(n - 1, parent, (const BvhNode*)c->nodes.p, (WideNode*)c->wnodes.p);
            c->wide_built = true;
            c->wide_compacted = false;
            if (c->wide_compact && n > 2) {
                if ((rc = dev_reserve(c, c->wnodes_c, nn * sizeof(WideNode))) || (rc = dev_reserve(c, c->wc_queue, nn * 16 + 8 * (size_t)c->num_sms * 16 + 64))) return rc;
                const WideNode* src = (const WideNode*)c->wnodes.p; WideNode* dst = (WideNode*)c->wnodes_c.p;
                int* qa = (int*)c->wc_queue.p; int* qb = qa + 2 * nn; int* ctl = qb + 2 * nn;       // ctl: 2 x grid CTA totals
                int per_sm = 0;
                PRT_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_compact_wide, 256, 0));
                const int grid = std::max(1, std::min((n + 255) / 256, c->num_sms * std::max(1, per_sm)));
                void* args[] = {&src, &dst, &qa, &qb, &ctl};
                PRT_CUDA(c, cudaLaunchCooperativeKernel((const void*)k_compact_wide, dim3(grid), dim3(256), args, 0, st));
                PRT_CUDA(c, cudaMemcpyAsync(&c->wide_live, ctl + 2 * grid, sizeof(int), cudaMemcpyDeviceToHost, st));   // read after the sync below
                c->wide_compacted = true;
            }
        }
        c->root_code = 0;
    } else {
        // single primitive: the root IS the leaf
        uint32_t payload = ns ? PRT_MAKE_LEAF(0, 0) : nq ? PRT_MAKE_LEAF(1, 0) : PRT_MAKE_LEAF(2, 0);
        c->root_code = ~(int)payload;
    }
    if (n == 1 && nt > 0)                                                 // n > 1: launched before the refit, above
        k_gather_tris<<<G, T, 0, st>>>(n, ns, nq, c->tri_prim_base, sprim, (const float4*)c->tri_in_v.p,
                                       (const float4*)c->tri_in_v.p + 3, (float4*)c->tri_v.p, (float4*)c->tri_n.p,
                                       (const float4*)c->aabb_min.p, (float4*)c->leaf_box.p, c->tri_deferred ? 1 : 0);
    if (c->tri_deferred && !c->finish_later && (rc = lbvh_finish_tris(c))) return rc;   // normals / materials into their leaf slots
    PRT_CUDA(c, cudaEventRecord(c->ev1, st));
    PRT_CUDA(c, cudaStreamSynchronize(st));
    PRT_CUDA(c, cudaGetLastError());
    PRT_CUDA(c, cudaEventElapsedTime(&c->build_ms, c->ev0, c->ev1));
    c->stats.build_ms = c->build_ms;
    c->bvh_built = true;
    return PRT_OK;
}

}  // namespace prt
