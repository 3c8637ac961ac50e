// prt_internal.h — context object and internal entry points shared by the .cu files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <utility>
#include <vector>
#include "../../include/prt.h"
#include "prt_device.cuh"

namespace prt {

struct DevBuf {
    void* p = nullptr; size_t bytes = 0;
};

// One captured 1-spp call (wavefront_render): everything the enqueue depends on except the first sample index and the seed.
struct GraphKey {
    uint64_t gen; int64_t pixel_begin, pixels; int max_depth, tail_b, counters; void* accum;
    bool operator==(const GraphKey& o) const {
        return gen == o.gen && pixel_begin == o.pixel_begin && pixels == o.pixels && max_depth == o.max_depth && tail_b == o.tail_b &&
               counters == o.counters && accum == o.accum;
    }
};
struct GraphEntry { GraphKey key; cudaGraphExec_t exec; int launches; };

struct Ctx {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    std::string err;

    // ---- scene (input order q: spheres, then quads, then triangles) ----
    int n_spheres = 0, n_quads = 0, n_tris = 0, n_prims = 0, n_mats = 0;
    int tri_prim_base = 0;
    DevBuf spheres, quads, quad_mat, mats;
    DevBuf raw_v, raw_n, raw_m;                 // the caller's packed triangle arrays as uploaded (kept: no allocation per prt_set_scene)
    void* stage[2] = {nullptr, nullptr};        // ring of two page-locked 32 MB buffers for pageable sources (upload_large)
    cudaEvent_t stage_ev[2] = {nullptr, nullptr};
    DevBuf tri_in_v, tri_in_n, tri_in_mat;     // tri_in_v, input order: 6 float4 / tri {v0, v1 (w = material), v2, n0, n1, n2}; the other two unused
    DevBuf prim_id_of_q;                        // int per q: reference list index
    float sky[3] = {0.4f, 0.3f, 0.6f};
    bool scene_set = false;
    // prt_set_scene_build: the input records hold positions only; normals / materials stay in raw_n / raw_m and every build ends
    // with k_finish_tris (tri_deferred). While the upload is still in flight the API layer runs that step itself (finish_later).
    bool tri_deferred = false, finish_later = false, def_normals = false, def_mat = false;
    int def_mat_default = 0;
    cudaStream_t up_stream = nullptr;           //   ... on this stream, from a second host thread

    // ---- LBVH ----
    DevBuf aabb_min, leaf_box;                  // {min, max} float4 pairs (32 B per box): per q in input order / per leaf slot in sorted order
    DevBuf scene_bounds;                        // 12 ordered-int floats: scene bounds, bounds of the box centres
    DevBuf tri_bounds;                          // the triangles' share of both, reduced when the scene is ingested
    DevBuf keys[2], vals[2];                    // radix sort ping-pong
    DevBuf block_hist;                          // digit-major per-block histograms
    DevBuf ray_keys[2][2], ray_vals[2][2], ray_hist[2];   // ray re-sorting, per buffer set: key/index ping-pong + histograms
    int morton_centroid = 1;                    // Morton grid over the bounds of the box centres (1) or the scene bounds (0)
    int sort_rays = 0;                          // staged pipeline: sort the queue of every bounce >= 1 (0 off, 1 on)
    int sort_key_bits = 24;                     // origin-cell Morton bits + 3 octant bits (8..24 => 1..3 radix passes)
    int sort_min_rays = 1 << 20;                // ... only while the queue is expected to hold at least this many rays
    DevBuf children, parent, node_min, node_max, flags;
    DevBuf nodes;                               // BvhNode per internal node
    DevBuf wnodes;                              // one WideNode per internal node (every node could be a wide root; prt_debug_wide returns this array)
    int bvh_wide = -1;                          // 4-wide quantised nodes for the staged extend: 1 always, 0 never, -1 by scene size
    int wide_min_prims = 4096;                  // auto: below this the two-child tree is L1-resident and its cheaper node step wins
    bool wide_built = false;
    int wide_compact = 1;                       // copy the reachable wide nodes (~45 %) into a dense, breadth-first array (lbvh.cu:k_compact_wide): soup 10 M +2.7 %, dragon 960 K +1.3 %, build +0.82 / +0.25 ms
    bool wide_compacted = false;
    int wide_live = 0;                          // records in the dense array
    DevBuf wnodes_c, wc_queue;
    DevBuf ploc_code, ploc_box, ploc_status, ploc_ctl;   // PLOC scratch: cluster codes (2 x n), boxes (the other half is node_min), tile status, control words
    int bvh_builder = 1;                        // hierarchy over the sorted leaves: 1 PLOC (falls back to Karras when it fails), 0 Karras (plain LBVH)
    int ploc_radius = 4;                        // neighbours searched on either side (1..8); 4 measured best (oracle counters: dragon, 809 spheres)
    int builder_used = 0, bvh_height = 0, ploc_rounds = 0;
    int ploc_trace[16 + 2 * 193] = {0};         // control words of the last PLOC build + (cluster count, ns timestamp) per round
    DevBuf tri_v, tri_n;                        // leaf-slot order (3 float4 per slot)
    DevBuf morton;                              // per q (kept for prt_debug_lbvh)
    int sorted_slot = 0;                        // which of keys/vals holds the sorted result
    int root_code = 0;
    bool bvh_built = false;
    float build_ms = 0.f;

    // ---- camera / film ----
    prt_camera cam{};
    bool cam_set = false;
    DevBuf accum_own;
    float* accum = nullptr;                     // W*H*3 (own or external)
    bool accum_external = false;
    int accum_spp = 0;
    DevBuf display;                             // RGBA8

    // ---- wavefront state ----
    int64_t wave_paths = 0;                     // capacity (paths in flight)
    int64_t wave_paths_opt = 0;                 // user override, 0 = auto
    DevBuf q_o[2], q_d[2], q_b[2];              // ray queues: float4 planes
    DevBuf hits;                                // float4 per ray
    DevBuf contrib;                             // float4 per path slot
    DevBuf qcount;                              // uint32 [PRT_MAX_DEPTH+2]
    // second buffer set + stream: consecutive waves alternate between the two sets so that the thin tail
    // bounces of wave k (a few thousand rays, launch-bound) run under the primary stage of wave k+1
    DevBuf q2_o[2], q2_d[2], q2_b[2], hits2, contrib2, qcount2;
    cudaStream_t stream2 = nullptr;
    cudaEvent_t ev_res[2] = {nullptr, nullptr}, ev_join = nullptr;
    int overlap = 1;                            // 1: two waves in flight (when a render has more than one wave)
    DevBuf stats_dev;                           // uint64 [PRT_MAX_DEPTH + 4]
    int ctas_per_sm = 8;
    int counters_on = 0;
    unsigned long long* stats_host = nullptr;   // pinned copy of stats_dev
    int tail_bounce = 2;               // fused path: bounces >= this run in ONE in-thread kernel, k_tail (0 = never); with
    int tail_adaptive = 1;             //   tail_adaptive the start is the first bounce whose queue is expected to hold <= tail_rays
    int tail_rays = 512 << 10;         //   rays (survival per bounce taken from the previous call on this scene; none yet: 1.0)
    int tail_rays_staged = 64 << 10;   //   same for the staged pipeline (its tail walks the two-child tree in-thread: hand over later)
    int tail_pool = 1;                 // fused scenes: the tail runs as a block-local wavefront (k_tail_pool) instead of one ray per thread
    int tail_max_paths = 1 << 26;      // waves with more paths never use the tail
    unsigned long long hint_rays[PRT_MAX_DEPTH] = {0};   // rays per bounce of the last fused render of this scene; [0] = 0: none yet
    int fused = -1;                             // 1: k_primary/k_bounce, 0: raygen/extend/shade as separate stages, -1: by scene size
    int fused_max_prims = 64;                   // auto: fuse when traversal is a handful of nodes (shading dominates)
    int brute_max_prims = 8;                    // fused kernels: sphere/quad scenes this small skip the BVH (uniform loop over the list)
    int persistent = 1;                         // staged extend = persistent threads with per-lane dynamic fetch
    int pt_refill_wide = 20;                    // same threshold for the 4-wide kernel (measured best 20: dragon960k, soup1m, soup10m)
    int pt_refill = 10;                         // refill the warp when fewer lanes than this are traversing (two-child walk, scenes below wide_min_prims; 809 spheres, Mrays/s: 4: 9 289, 8: 9 580, 10: 9 593, 12: 9 554, 14: 9 479, 16: 9 360)
    int rr_start = -1;                          // estimator options, both OFF by default so that parity with the
    float firefly_clamp = 0.f;                  // reference's estimator stays checkable: Russian roulette / radiance clamp
    int stage_timing = 1;                       // CUDA-event pairs around the hot kernels
    int fast_math = 0;                          // option: SFU approximations for divide / sqrt / rsqrt on the SHADING side (default: bit-exact IEEE)
    int use_graph = 1;                          // one-sample calls replay a captured CUDA graph (option "use_graph")
    uint64_t graph_gen = 0;                     // bumped by everything that changes what a call enqueues (scene, BVH, camera, options, buffers)
    std::vector<GraphEntry> graphs;
    uint32_t* dyn_host = nullptr;               // pinned {first sample, seed_lo, seed_hi}: read by a memcpy node at replay time
    DevBuf dyn_dev;
    std::vector<cudaEvent_t> ev_pool;
    std::vector<std::pair<size_t, int>> pend_spans;   // a render that is enqueued but not finished yet (wavefront_finish)
    int pend_launches = 0; uint64_t pend_paths = 0; bool pend_active = false;
    prt_stats stats{};

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

// error helpers ------------------------------------------------------------------
int set_err(Ctx* c, int code, const std::string& msg);
#define PRT_CUDA(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return prt::set_err((ctx), e_ == cudaErrorMemoryAllocation ? PRT_ERR_OOM : PRT_ERR_CUDA, \
        std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

int dev_reserve(Ctx* c, DevBuf& b, size_t bytes);   // grow-only allocation
void dev_free(DevBuf& b);

SceneView scene_view(const Ctx* c);

// lbvh.cu
int lbvh_build(Ctx* c);
int lbvh_ingest_tris(Ctx* c, int n_spheres, int n_quads, int n_tris, bool have_normals, bool have_mat, int mat_default, bool positions_only = false);
int lbvh_finish_tris(Ctx* c);
int sort_ray_queue(Ctx* c, cudaStream_t st, int set, int64_t capacity, const uint32_t* n_dev, const float4* so, const float4* sd,
                   const float4* sb, float4* to, float4* td, float4* tb);
// wavefront.cu
int wavefront_render(Ctx* c, int spp, int max_depth, uint64_t seed, int sample_begin, int64_t pixel_begin, int64_t pixel_count, bool fresh = false,
                     bool defer_finish = false);
int wavefront_finish(Ctx* c);
void wavefront_drop_graphs(Ctx* c);     // forget every captured call (anything it depends on changed)
unsigned wavefront_check_flags();            // checked builds (-DPRT_CHECKED=1): violated-index bits since the last call, 0 otherwise
unsigned lbvh_check_flags();
void wavefront_set_stack_limit(int limit);
int wavefront_trace_batch(Ctx* c, const prt_ray* rays, int n, prt_hit* out, int use_bvh);
int film_resolve_display(Ctx* c, float exposure, float gamma, uint8_t* host_rgba);
int wavefront_debug_scatter(Ctx* c, const prt_material* mats, const float* in_dirs, const float* normals, const int32_t* fronts,
                            const uint32_t* keys, uint64_t seed, int n, float* out);
int film_scale_accum(Ctx* c, float s);

}  // namespace prt

// A context handle. `c` is the context of the one device (prt_create) or of the ROOT device (prt_create_multi);
// a multi-device context owns one more handle per additional device.
struct prt_ctx {
    prt::Ctx c;
    std::vector<prt_ctx*> peers;            // multi-device: contexts of device_ids[1..]
    std::vector<void*> comms;               // ncclComm_t per device (root first); empty = NCCL not in use
    int reduce_mode = 0;                    // 0 auto (NCCL when libnccl.so.2 loads, else peer-to-peer kernel), 1 NCCL, 2 peer-to-peer
    int partition = PRT_PARTITION_SAMPLES;
    bool p2p_enabled = false;
    cudaEvent_t evr0 = nullptr, evr1 = nullptr;
    std::vector<cudaEvent_t> ev_peer;       // peer-to-peer reduce: "this device's waves are done", one per peer
};
