/*
 * prt.h — C ABI of the B200-native path-tracing backend (libprt_b200.so).
 *
 * This is the drop-in boundary for ONE hot path of Rickyeeeeee/ParallelRayTracing:
 *   camera ray -> closest hit -> shade/scatter -> accumulate
 * i.e. what a `Renderer` backend does between `Init` and the hand-over of one
 * sample buffer to `Film` (reference: src/core/renderer.h:8-16).
 *
 * Plain pointers and sizes only; no C++/torch types.  Every entry point returns an
 * int status (0 = PRT_OK, <0 = PRT_ERR_*) and never throws.  A context is NOT
 * thread-safe; it is driven from one host thread, like the reference's backends
 * (main.cpp:194-549 calls everything from the GL thread).
 *
 * All citations are relative to the reference tree (/root/reference).
 */
#ifndef PRT_H
#define PRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PRT_OK                0
#define PRT_ERR_INVALID      -1  /* bad argument / call order            */
#define PRT_ERR_CUDA         -2  /* a CUDA runtime call failed            */
#define PRT_ERR_NO_DEVICE    -3  /* no usable sm_100 device               */
#define PRT_ERR_OOM          -4  /* device allocation failed              */
#define PRT_ERR_UNSUPPORTED  -5

/* Material tags: identical numbering to MatType (src/core/material_handle.h:13-19). */
#define PRT_MAT_NONE        0
#define PRT_MAT_LAMBERTIAN  1
#define PRT_MAT_METAL       2
#define PRT_MAT_DIELECTRIC  3
#define PRT_MAT_EMISSIVE    4

/* Shape tags: identical numbering to ShapeType (src/core/shape.h:10-15). */
#define PRT_SHAPE_SPHERE    0   /* "Circle" in the reference */
#define PRT_SHAPE_QUAD      1
#define PRT_SHAPE_TRIANGLE  2

/* How a multi-device context (prt_create_multi) splits one prt_render call over its devices; option
 * "multi_partition". Either way every device renders into its own buffer and ONE sum-reduce lands the frame on device 0. */
#define PRT_PARTITION_SAMPLES 0  /* device d renders its share of the sample range [sample_begin, sample_begin+spp) of every pixel */
#define PRT_PARTITION_TILES   1  /* device d renders all samples of its band of rows                                             */

typedef struct prt_ctx prt_ctx;

/* One material.  Replaces the 4 material structs + MaterialHandle tag dispatch
 * (src/core/material.h:11-161).  rgb = albedo (Lambertian, Metal) or emission
 * (Emissive); param = roughness (Metal) or refraction index (Dielectric). */
typedef struct prt_material {
    int32_t type;
    float   r, g, b;
    float   param;
    int32_t _pad[3];
} prt_material;              /* 32 B */

/* World-space sphere = reference `Circle` (shape.h:157-203) with its Transform baked in
 * (uniform scale s): centre = Mat*(0,0,0), radius = r*s, tmin = kShapeRayTMin*s
 * (shape.h:128 is in LOCAL units, primitive.cpp:30 re-normalises the local direction). */
typedef struct prt_sphere {
    float   cx, cy, cz, radius;
    float   tmin;
    int32_t material;        /* index into the material table                      */
    int32_t prim_id;         /* index in the reference's primitive list (tie rule) */
    int32_t _pad;
} prt_sphere;                /* 32 B */

/* World-space rectangle = reference `Quad` (shape.h:213-239; local y=0 plane, extent
 * width x height along local x / z) with its Transform baked in. u, v, n are the unit
 * images of local x, z, y. */
typedef struct prt_quad {
    float   cx, cy, cz, half_w;
    float   ux, uy, uz, half_h;
    float   vx, vy, vz, tmin;
    float   nx, ny, nz;
    int32_t material;
    int32_t prim_id;
    int32_t _pad[3];
} prt_quad;                  /* 80 B */

/* Camera = the fields Camera::GetCameraRay reads (src/core/camera.h:104-143).
 * right = normalize(cross(front, (0,1,0))), up = normalize(cross(right, front))
 * (camera.h:13-15); tan_half_fov = tan(0.5f) in the reference (camera.h:111). */
typedef struct prt_camera {
    float   pos[3];
    float   front[3];
    float   right[3];
    float   up[3];
    float   tan_half_fov;
    int32_t width, height;
} prt_camera;

/* A ray of the fixed-batch parity interface. Direction must be unit length. */
typedef struct prt_ray { float ox, oy, oz, dx, dy, dz; } prt_ray;

/* Closest-hit record: prim_id = reference list index (or -1 = miss), t = world distance,
 * (b1, b2) = triangle barycentrics (weights of v1, v2), front = IsFrontFace. */
typedef struct prt_hit {
    float   t;
    int32_t prim_id;
    float   b1, b2;
    float   px, py, pz;      /* SurfaceInteraction::Position */
    float   nx, ny, nz;      /* SurfaceInteraction::Normal (faces the ray) */
    int32_t front;
    int32_t material;
} prt_hit;                   /* 48 B */

#define PRT_MAX_DEPTH 32

typedef struct prt_stats {
    uint64_t rays_total;                  /* closest-hit queries of the last prt_render      */
    uint64_t rays_per_bounce[PRT_MAX_DEPTH];
    uint64_t paths;                       /* camera paths started                              */
    uint64_t nodes_visited;               /* BVH nodes fetched   (only if counters enabled)    */
    uint64_t prims_tested;                /* leaf primitives tested (only if counters enabled) */
    float    render_ms;                   /* device time of the last prt_render (CUDA events)  */
    float    extend_ms;                   /* device time in the extend kernels (fused mode: k_primary + k_bounce) */
    float    shade_ms;
    float    build_ms;                    /* device time of the last prt_build_bvh             */
    float    primary_ms;                  /* part of extend_ms spent in the bounce-0 kernels   */
    float    reduce_ms;                   /* multi-device context: device time of the sum-reduce (incl. waiting for the slowest device) */
    int32_t  kernel_launches;             /* kernels launched by the last prt_render           */
    int32_t  n_prims;
    int32_t  n_nodes;
    int32_t  primary_launches;            /* bounce-0 kernel launches (= waves) of the last prt_render */
    int32_t  bvh_builder;                 /* hierarchy of the last prt_build_bvh: 1 PLOC, 0 Karras (LBVH; also the fallback) */
    int32_t  bvh_height;                  /* height of the PLOC tree (0 for Karras: bounded by the key width instead)     */
    int32_t  ploc_rounds;                 /* clustering rounds the PLOC build took                                         */
    int32_t  n_devices;                   /* devices of this context (prt_create: 1)                                        */
} prt_stats;

/* ---- life cycle --------------------------------------------------------------------
 * Replaces a backend's ctor/dtor + the allocation half of Init
 * (cuda_wavefront/renderer.cu:351-370, 436-487). device = CUDA ordinal. */
int  prt_create(prt_ctx** out, int device);
/* ONE context over n_devices GPUs of this process (SURVEY 8b export list / 8e: single process, one stream per device,
 * ncclCommInitAll). device_ids[0] is the root: its accumulation buffer is the one every other entry point reads. The scene,
 * the BVH (built on every device, identical bits) and the camera are replicated; prt_render* split the call over the devices
 * (option "multi_partition": PRT_PARTITION_SAMPLES default, PRT_PARTITION_TILES), enqueue every device's waves and, behind
 * them on the same streams with no host synchronisation in between, ONE float sum-reduce to the root: ncclReduce over
 * NVLink (libnccl.so.2 is loaded at run time; option "multi_reduce" = 1), or a peer-to-peer kernel on the root that reads
 * the other devices' buffers through NVLink peer mappings (= 2; also the fallback when NCCL cannot be loaded; 0 = auto).
 * The image equals the one-device image up to fp32 summation order (same samples: the RNG is keyed by pixel and sample). */
int  prt_create_multi(prt_ctx** out, const int* device_ids, int n_devices);
void prt_destroy(prt_ctx* ctx);
const char* prt_last_error(const prt_ctx* ctx);   /* ctx may be NULL: last global error */
const char* prt_version(void);

/* ---- scene -------------------------------------------------------------------------
 * Replaces the scene-upload half of Init (cuda_wavefront/soa.cpp:37-114,
 * cuda_megakernel/renderer.cu:276-355): the host scene is COPIED here, host pointers are
 * never dereferenced on the device.
 * Triangles = reference `Triangle` (shape.h:49-82) in world space: tri_verts = nt*9 floats
 * (v0,v1,v2), tri_normals = nt*9 floats or NULL (flat geometric normals), tri_material =
 * nt ints or NULL (all tri_material_default). Triangle i has prim_id tri_prim_base + i.
 * sky_rgb = radiance of a miss (the reference hard-codes (0.4,0.3,0.6): cpu/renderer.h:31).
 * Non-finite input: primitives with NaN / +-inf coordinates are tolerated - they are never hit and do not disturb the
 * build or the hits on the others (tests/test_checked_build.py). Finite coordinates whose products overflow fp32
 * (|x| beyond ~1e18) give unspecified hits at t = inf, as they do in the reference's own tests. */
int  prt_set_scene(prt_ctx* ctx,
                   const prt_sphere* spheres, int n_spheres,
                   const prt_quad* quads, int n_quads,
                   const float* tri_verts, const float* tri_normals,
                   const int32_t* tri_material, int tri_material_default,
                   int n_tris, int tri_prim_base,
                   const prt_material* materials, int n_materials,
                   const float sky_rgb[3]);

/* prt_set_scene + prt_build_bvh in one call - what a backend's Init does (cuda_wavefront/renderer.cu:351-370 uploads the scene and
 * allocates in one go; there is no build in the reference). Same arguments, same device state afterwards, bit for bit; the
 * difference is time: the hierarchy is built from the vertex positions (boxes, Morton codes, sort, PLOC, wide nodes, vertex
 * gather) as soon as THEY are on the device, while a second host thread is still sending the vertex normals and material
 * indices - 40 of the 76 bytes per triangle - over PCIe; a last kernel drops those into their leaf slots. The BVH build of a
 * 10 M-triangle mesh (3.8 ms) disappears behind 7 ms of upload. Synchronous: the arrays are free again on return. */
int  prt_set_scene_build(prt_ctx* ctx,
                         const prt_sphere* spheres, int n_spheres,
                         const prt_quad* quads, int n_quads,
                         const float* tri_verts, const float* tri_normals,
                         const int32_t* tri_material, int tri_material_default,
                         int n_tris, int tri_prim_base,
                         const prt_material* materials, int n_materials,
                         const float sky_rgb[3]);

/* Page-locked host memory (cudaHostAlloc) for the big input arrays: prt_set_scene copies from such memory with plain
 * asynchronous DMA (~50 GB/s over PCIe 5); pageable arrays go through the context's own staging ring instead. Optional. */
void* prt_host_alloc(size_t bytes);
void  prt_host_free(void* p);

/* GPU LBVH build over the current scene: primitive AABBs -> scene bounds -> 30-bit Morton
 * codes -> stable LSD radix sort -> hierarchy over the sorted leaves: PLOC (locally-ordered agglomerative clustering,
 * the tree-quality pass of SURVEY 8f-4; option bvh_builder=1, default) or the Karras hierarchy + bottom-up refit
 * (bvh_builder=0; also the automatic fallback when PLOC does not converge or its tree is too deep for the stack).
 * The reference has no acceleration structure (primitive.cpp:26 loops over every
 * primitive); this replaces that loop. */
int  prt_build_bvh(prt_ctx* ctx);

/* Replaces SetCamera (cuda_wavefront/renderer.cu:534-547). Resizes the accumulation
 * buffer (and clears it) when width/height change. */
int  prt_set_camera(prt_ctx* ctx, const prt_camera* cam);

/* ---- render ------------------------------------------------------------------------
 * Replaces ProgressiveRender (cuda_wavefront/renderer.cu:372-433), batched: renders
 * samples [sample_begin, sample_begin + spp) of every pixel (pixel centre, no jitter:
 * cpu/renderer.cpp:45) with at most max_depth closest-hit queries per path
 * (cpu/renderer.cpp:59-103 semantics) and ADDS them to the accumulation buffer.
 * Synchronous, like the reference (cudaDeviceSynchronize at renderer.cu:429). */
int  prt_render(prt_ctx* ctx, int spp, int max_depth, uint64_t seed, int sample_begin);
/* "Clear, then render" in one call: afterwards the sample buffer holds exactly the sum of these `spp` samples and
 * prt_accum_spp() == spp. This is what one Renderer::ProgressiveRender() needs before it hands the buffer to
 * Film::AddSampleBufferGPU (cuda_wavefront/renderer.cu:372-433 refills its own m_DeviceFilmBuffer every frame the same
 * way). With spp = 1 and a frame that fits one wave the paths write their contribution straight into the buffer: no
 * memset, no contribution plane, no resolve pass. Same image, bit for bit, as prt_clear_accum + prt_render. */
int  prt_render_fresh(prt_ctx* ctx, int spp, int max_depth, uint64_t seed, int sample_begin);

/* The same for the pixel range [pixel_begin, pixel_begin + pixel_count) only (row-major pixel index y*W+x; a band
 * of rows is a contiguous range): the TILE partition of SURVEY 8e - each GPU renders all samples of its own pixels,
 * pixels outside the range are left untouched, and the same sum-reduce assembles the frame. Identical per-pixel
 * results to prt_render (the RNG is keyed by pixel and sample, not by launch geometry). Does not change the
 * weight reported by prt_accum_spp unless the range is the whole frame. */
int  prt_render_region(prt_ctx* ctx, int spp, int max_depth, uint64_t seed, int sample_begin, int pixel_begin, int pixel_count);

/* Make the context's render stream wait (on the device, the host does not block) for a CUDA event of the caller's -
 * cuda_event is a cudaEvent_t - before anything enqueued afterwards runs. The Renderer shim records such an event behind
 * Film::AddSampleBufferGPU (which reads the sample buffer on the legacy default stream and returns without waiting,
 * film.cu:90-99), so that the next ProgressiveRender cannot overwrite the buffer early - without a host synchronisation. */
int  prt_wait_event(prt_ctx* ctx, void* cuda_event);

/* Zero the accumulation buffer (Film::Clear, film.cu:25-35). */
int  prt_clear_accum(prt_ctx* ctx);

/* Multiply the accumulation buffer by s (used by the Renderer shim to hand Film a mean
 * with weight k: Film::AddSampleBufferGPU adds rgb*weight, film.cu:79-88). */
int  prt_scale_accum(prt_ctx* ctx, float s);

/* Device pointer to the accumulation buffer: W*H*3 fp32, interleaved RGB, row-major,
 * index 3*(y*W+x)+c — the layout Film::AddSampleBufferGPU consumes (film.h:26-29,
 * film.cu:79-88). Holds the SUM of all samples since the last clear. */
float* prt_accum_device_ptr(prt_ctx* ctx);
/* Use caller-owned device memory (W*H*3 floats) as the accumulation buffer, e.g. a
 * torch tensor that an NCCL reduce then operates on. NULL restores the internal one. */
int  prt_set_accum_buffer(prt_ctx* ctx, float* device_rgb);
/* Copy the accumulation buffer to host memory (W*H*3 floats). */
int  prt_read_accum(prt_ctx* ctx, float* host_rgb);
/* Samples per pixel accumulated since the last clear. */
int  prt_accum_spp(const prt_ctx* ctx);
/* Declare the weight (samples per pixel) of what the accumulation buffer holds. For a caller that fills
 * the external buffer itself - e.g. rank 0 after the NCCL sum-reduce of N ranks' buffers - so that
 * prt_resolve_display divides by the right count (Film::m_Weights after AddSampleBufferGPU(buf, w),
 * film.cu:79-99, 111-115). */
int  prt_set_accum_spp(prt_ctx* ctx, int spp);

/* ---- film resolve -----------------------------------------------
 * accum/spp -> Reinhard -> gamma -> RGBA8, Film::UpdateDisplay semantics
 * (film.cu:134-194). host_rgba: W*H*4 bytes. */
int  prt_resolve_display(prt_ctx* ctx, float exposure, float gamma, uint8_t* host_rgba);

/* ---- parity / introspection --------------------------------------------------------
 * Closest hit for a fixed ray batch (host buffers). use_bvh=0 brute-forces all
 * primitives on the GPU (the reference's own algorithm, primitive.cpp:21-59); 1 walks the
 * two-child LBVH (node/leaf counters comparable with the CPU oracle); 2 walks the 4-wide
 * quantised nodes the render path uses (same hits, fewer node visits). */
int  prt_trace_batch(prt_ctx* ctx, const prt_ray* rays, int n, prt_hit* out, int use_bvh);

/* prt_trace_batch for callers that cannot allocate (stream capture, tight serving loops, harnesses that forbid cudaMalloc).
 * NO COUNTERPART IN THE REFERENCE. Same kernels and the same hit records, bit
 * for bit, as prt_trace_batch, but
 *   - rays_dev / out_dev are DEVICE-accessible pointers of the caller's (device memory of the context's device, or pinned /
 *     managed memory the GPU can address); the kernels read and write them directly, nothing is copied;
 *   - the scratch (ray queue, hit plane, counters) is a static device array of the library, sized for at most
 *     PRT_TRACE_STATIC_MAX_RAYS rays (3 MB + 1 KB, held by every GPU that loads the library until the process exits);
 *     n above that returns PRT_ERR_INVALID with a message, n = 0 only resolves the array's address (the once-per-device set-up);
 *   - no cudaMalloc / cudaFree / allocating routine is called, prt_stats is not updated;
 *   - cuda_stream is the cudaStream_t to enqueue on (NULL = the context's render stream). The call waits for that stream
 *     before it returns, unless the stream is being captured, in which case the chain is only recorded.
 * Every call on one GPU shares the one array: calls that overlap on the GPU (two streams, two contexts, two graph launches
 * in flight) race on it - serialise them. A multi-device context traces on its root device. */
#define PRT_TRACE_STATIC_MAX_RAYS 65536
int  prt_trace_batch_static(prt_ctx* ctx, const prt_ray* rays_dev, int n, prt_hit* out_dev, int use_bvh, void* cuda_stream);

/* MaterialHandle::{Emit,Scatter} known answers (src/core/material.h:11-161, math.h:10-50) through the device function the
 * render path shades with. Record k: material mats[k], incoming unit direction in_dirs[3k..], hit normal facing the ray
 * normals[3k..], IsFrontFace fronts[k], RNG stream of (seed, pixel = keys[3k], sample = keys[3k+1], bounce = keys[3k+2]).
 * out: n x 10 floats {scattered (0/1), attenuation rgb, scattered direction (normalised; 0 when the path ends), emission rgb}.
 * Host buffers. The reference's own answers for 600 such records are in tests/golden (made by oracle/ref_harness.cpp). */
int  prt_debug_scatter(prt_ctx* ctx, const prt_material* mats, const float* in_dirs, const float* normals,
                       const int32_t* fronts, const uint32_t* keys, uint64_t seed, int n, float* out);

/* LBVH internals for the bit-exact integer checks. Any pointer may be NULL.
 *   morton[n_prims]        30-bit code per primitive, input order
 *   sorted_prim[n_prims]   primitive index (input order) at each sorted position
 *   node_children[2*(n-1)] left,right per internal node; >=0 internal, <0 = ~leaf_pos
 *   node_parent[n-1]       parent of each internal node (root: -1)
 *   node_bounds[12*(n-1)]  left box min,max then right box min,max per internal node
 *   scene_bounds[6]        min,max of all primitive AABBs */
int  prt_debug_lbvh(prt_ctx* ctx, uint32_t* morton, uint32_t* sorted_prim,
                    int32_t* node_children, int32_t* node_parent,
                    float* node_bounds, float* scene_bounds);
/* The 4-wide quantised nodes (only when built: option bvh_wide, or automatically for large scenes), raw device
 * layout: 16 words per internal node {ox, oy, oz, scale_x, code[4], qlo[3], qhi[3], scale_y, scale_z}, indexed like
 * the two-child nodes; every internal node has a record (greedy collapse by surface area, 2..4 slots). The walk itself
 * reads a dense breadth-first copy of the records reachable from the root (option wide_compact, default on; same visits,
 * hits and counters). Returns PRT_ERR_INVALID if not built. */
int  prt_debug_wide(prt_ctx* ctx, uint32_t* words);
/* The dense copy the walk reads (option wide_compact, default on): the records reachable from the root in breadth-first order,
 * root first, a node's internal children at consecutive indices, internal child codes = indices into THIS array. *n_nodes =
 * number of records; words (capacity_nodes x 16 words) may be NULL to ask for the count only. Deterministic: the same
 * bytes on every run and device (tests rebuild it from prt_debug_wide's array on the host and compare). */
int  prt_debug_wide_compact(prt_ctx* ctx, uint32_t* words, int capacity_nodes, int* n_nodes);
/* The repo's own memcheck (compute-sanitizer is closed on the GPU pool): libprt_b200_checked.so is the same library compiled
 * with -DPRT_CHECKED=1 - every data-derived index in device code (traversal stack pointers, queue slots, node and leaf
 * codes, PLOC positions and node ids, sort scatter positions) is tested before the access, violations set PRT_CHK_* bits
 * (prt_device.cuh) and skip the access. Returns the bits raised since the last call, OR-ed with 1<<30 (= "this is the
 * checked library"); the regular library returns 0. */
int  prt_debug_check_flags(prt_ctx* ctx);

/* Round trace of the last PLOC build: counts[r] = clusters at the start of round r, ns[r] = device %globaltimer (low 32 bits,
 * nanoseconds) at that moment; returns the number of entries written (rounds + 1, <= max_rounds) or < 0. Profiling aid. */
int  prt_debug_ploc_trace(prt_ctx* ctx, int32_t* counts, uint32_t* ns, int max_rounds);
int  prt_num_prims(const prt_ctx* ctx);

/* Enable per-ray traversal counters (nodes fetched / primitives tested); costs speed. */
int  prt_enable_counters(prt_ctx* ctx, int on);
int  prt_get_stats(prt_ctx* ctx, prt_stats* out);

/* Tuning knobs (all optional): "wave_paths" (paths in flight per wave), "ctas_per_sm", "fused", "fused_max_prims",
 * "brute_max_prims", "persistent", "pt_refill", "pt_refill_wide", "bvh_wide", "wide_min_prims", "wide_compact", "bvh_builder", "ploc_radius", "overlap",
 * "stage_timing", "use_graph", "fast_math" (tolerance-mode float math on the shading side: changes radiance within RMSE 1e-3 / max-rel 1e-2, never a hit id), "rr_start", "firefly_clamp_milli", "multi_partition", "multi_reduce" (multi-device contexts), and for small waves (one sample per call, the reference's
 * ProgressiveRender granularity) "tail_bounce" (0 = one launch per bounce always), "tail_adaptive", "tail_pool", "tail_rays", "tail_rays_staged",
 * "tail_max_paths": bounces from the tail start on run inside one kernel; "sort_rays", "sort_key_bits",
 * "sort_min_rays": re-sort the ray queue of every bounce by origin cell and direction octant (staged pipeline, off). None of them changes the image except
 * rr_start / firefly_clamp_milli (estimator options, default off). Unknown keys -> PRT_ERR_INVALID. */
int  prt_set_option(prt_ctx* ctx, const char* key, int64_t value);

#ifdef __cplusplus
}
#endif
#endif /* PRT_H */
