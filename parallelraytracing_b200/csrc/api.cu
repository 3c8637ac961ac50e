// api.cu — the extern "C" boundary declared in include/prt.h.
// Replaces the host half of a reference backend: Init's allocation/upload
// (src/backend/cuda_wavefront/renderer.cu:351-370, soa.cpp:37-114), SetCamera (:534-547)
// and ProgressiveRender's driver loop (:372-433).
#include "prt_internal.h"
#include <cstring>
#include <cstdio>
#include <algorithm>
#include <thread>
#include <dlfcn.h>
#include <nccl.h>      // types and enums only: every NCCL symbol is resolved with dlsym, the library has no link-time dependency on it

namespace prt {

static thread_local std::string g_err;

int set_err(Ctx* c, int code, const std::string& msg) {
    g_err = msg;
    if (c) c->err = msg;
    return code;
}

int dev_reserve(Ctx* c, DevBuf& b, size_t bytes) {
    if (bytes == 0) bytes = 16;
    if (b.p && b.bytes >= bytes) return PRT_OK;
    if (b.p) { cudaFree(b.p); b.p = nullptr; b.bytes = 0; }
    cudaError_t e = cudaMalloc(&b.p, bytes);
    if (e != cudaSuccess) {
        b.p = nullptr;
        cudaGetLastError();
        char buf[160];
        snprintf(buf, sizeof buf, "cudaMalloc(%zu bytes): %s", bytes, cudaGetErrorString(e));
        return set_err(c, PRT_ERR_OOM, buf);
    }
    b.bytes = bytes;
    return PRT_OK;
}

void dev_free(DevBuf& b) { if (b.p) cudaFree(b.p); b.p = nullptr; b.bytes = 0; }

static int upload(Ctx* c, DevBuf& b, const void* host, size_t bytes) {
    int rc = dev_reserve(c, b, bytes);
    if (rc) return rc;
    if (bytes) PRT_CUDA(c, cudaMemcpyAsync(b.p, host, bytes, cudaMemcpyHostToDevice, c->stream));
    return PRT_OK;
}

// Large host arrays (the packed triangle arrays: 76 bytes per triangle, 760 MB for 10 M). Page-locked source (the caller
// used prt_host_alloc / cudaHostAlloc / cudaHostRegister): one asynchronous DMA per 32 MB chunk straight from the caller's
// memory. Pageable source: the driver's own staging path did 5-9 GB/s here (88-161 ms for the 10 M soup), so
// the copy goes through the context's ring of two page-locked 32 MB buffers instead - four host threads fill one buffer
// while the DMA engine drains the other (one cudaMemcpyAsync per chunk).
static void parallel_memcpy(void* dst, const void* src, size_t bytes, int n_threads) {
    if (bytes < (4u << 20) || n_threads <= 1) { memcpy(dst, src, bytes); return; }
    std::vector<std::thread> th;
    const size_t per = (bytes / n_threads + 4095) & ~(size_t)4095;
    for (int t = 1; t < n_threads; ++t) {
        const size_t off = (size_t)t * per;
        if (off >= bytes) break;
        th.emplace_back([=] { memcpy((char*)dst + off, (const char*)src + off, std::min(per, bytes - off)); });
    }
    memcpy(dst, src, std::min(per, bytes));
    for (auto& t : th) t.join();
}

static int upload_large(Ctx* c, DevBuf& b, const void* host, size_t bytes, cudaStream_t st = nullptr, bool reserve = true) {
    if (!st) st = c->stream;
    int rc = reserve ? dev_reserve(c, b, bytes) : PRT_OK;
    if (rc || !bytes) return rc;
    const size_t CH = (size_t)32 << 20;
    cudaPointerAttributes at{};
    const bool pinned = cudaPointerGetAttributes(&at, host) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    if (pinned || bytes < ((size_t)8 << 20)) {
        for (size_t off = 0; off < bytes; off += CH)
            PRT_CUDA(c, cudaMemcpyAsync((char*)b.p + off, (const char*)host + off, std::min(CH, bytes - off), cudaMemcpyHostToDevice, st));
        return PRT_OK;
    }
    for (int k = 0; k < 2; ++k) {
        if (!c->stage[k]) PRT_CUDA(c, cudaHostAlloc(&c->stage[k], CH, cudaHostAllocDefault));
        if (!c->stage_ev[k]) PRT_CUDA(c, cudaEventCreateWithFlags(&c->stage_ev[k], cudaEventDisableTiming));
    }
    int k = 0;
    for (size_t off = 0; off < bytes; off += CH, k ^= 1) {
        const size_t n = std::min(CH, bytes - off);
        PRT_CUDA(c, cudaEventSynchronize(c->stage_ev[k]));             // the DMA that last read this buffer is done (no-op the first time)
        parallel_memcpy(c->stage[k], (const char*)host + off, n, 4);
        PRT_CUDA(c, cudaMemcpyAsync((char*)b.p + off, c->stage[k], n, cudaMemcpyHostToDevice, st));
        PRT_CUDA(c, cudaEventRecord(c->stage_ev[k], st));
    }
    return PRT_OK;
}

// ------------------------------------------------------------------ multi-device reduce
// NCCL entry points, loaded once per process from libnccl.so.2 (the copy torch already mapped when there is one).
struct NcclApi {
    bool tried = false, ok = false;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool load() {
        if (tried) return ok;
        tried = true;
        void* lib = nullptr;
        const char* env = getenv("PRT_NCCL_LIB");
        if (env && *env) lib = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) return false;
        CommInitAll = (decltype(CommInitAll))dlsym(lib, "ncclCommInitAll");
        CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
        GroupStart = (decltype(GroupStart))dlsym(lib, "ncclGroupStart");
        GroupEnd = (decltype(GroupEnd))dlsym(lib, "ncclGroupEnd");
        Reduce = (decltype(Reduce))dlsym(lib, "ncclReduce");
        GetErrorString = (decltype(GetErrorString))dlsym(lib, "ncclGetErrorString");
        ok = CommInitAll && CommDestroy && GroupStart && GroupEnd && Reduce && GetErrorString;
        return ok;
    }
};
static NcclApi g_nccl;

// Peer-to-peer alternative: the root adds the other devices' buffers to its own, reading them through NVLink peer
// mappings (one kernel, 16-byte loads; n_peers <= 15 pointers travel by value).
struct PeerPtrs { const float* p[15]; };
__global__ void k_sum_peers(float* __restrict__ dst, PeerPtrs src, int n_peers, size_t n) {
    const size_t n4 = n / 4, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 a = reinterpret_cast<float4*>(dst)[i];
        for (int k = 0; k < n_peers; ++k) {
            const float4 b = __ldcs(reinterpret_cast<const float4*>(src.p[k]) + i);
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        reinterpret_cast<float4*>(dst)[i] = a;
    }
    if (blockIdx.x == 0 && threadIdx.x < (unsigned)(n - 4 * n4)) {
        const size_t i = 4 * n4 + threadIdx.x;
        float a = dst[i];
        for (int k = 0; k < n_peers; ++k) a += src.p[k][i];
        dst[i] = a;
    }
}

}  // namespace prt

using namespace prt;

// every device of a handle, root first
static std::vector<prt_ctx*> all_devices(prt_ctx* h) {
    std::vector<prt_ctx*> v{h};
    v.insert(v.end(), h->peers.begin(), h->peers.end());
    return v;
}

static bool argcheck_render(Ctx* c, const char* who, int spp, int max_depth, int sample_begin, int min_spp) {
    if (!c->scene_set || !c->bvh_built) { set_err(c, PRT_ERR_INVALID, std::string(who) + ": scene/BVH not ready (prt_set_scene + prt_build_bvh)"); return false; }
    if (!c->cam_set) { set_err(c, PRT_ERR_INVALID, std::string(who) + ": call prt_set_camera first"); return false; }
    if (spp < min_spp || max_depth < 0 || max_depth > PRT_MAX_DEPTH || sample_begin < 0) {
        set_err(c, PRT_ERR_INVALID, std::string(who) + ": bad spp / max_depth / sample_begin"); return false;
    }
    return true;
}

// One render call of a multi-device context: split, enqueue everywhere, reduce behind the waves, then wait.
//   samples: device d renders samples [b_d, b_d + n_d) of the pixel range; tiles: all samples of its share of the pixel range
//   (whole rows when the range is the whole frame). The root ADDS into its accumulation buffer (or clears it first: `fresh`),
//   every other device clears its own buffer and renders into it; the sum-reduce then adds those to the root's.
static int multi_render(prt_ctx* h, int spp, int max_depth, uint64_t seed, int sample_begin, int64_t pixel_begin, int64_t pixel_count, bool fresh) {
    std::vector<prt_ctx*> dev = all_devices(h);
    const int N = (int)dev.size();
    Ctx* root = &h->c;
    const int W = root->cam.width;
    const size_t n_floats = (size_t)W * root->cam.height * 3;
    const int base_spp = fresh ? 0 : root->accum_spp;
    const bool tiles = h->partition == PRT_PARTITION_TILES;
    int rc = PRT_OK;
    // One host thread per device enqueues that device's waves: a long render is thousands of launches and the driver's launch
    // queue is finite, so a single thread would block inside device 0's launches until device 0 is nearly done and only then
    // start feeding device 1 (measured: 2 GPUs, 4K x 1024 spp: device 1 idle for 288 ms of 1363).
    std::vector<int> rcs(N, PRT_OK);
    auto enqueue = [&](int d) {
        Ctx* c = &dev[d]->c;
        if (cudaSetDevice(c->device) != cudaSuccess) { rcs[d] = set_err(c, PRT_ERR_CUDA, "cudaSetDevice failed"); return; }
        int n_d = spp, b_d = sample_begin; int64_t p_d = pixel_begin, pc_d = pixel_count;
        if (tiles) {                                   // contiguous pixel shares, cut at row boundaries when the range starts on one
            const int64_t unit = (pixel_begin % W == 0 && pixel_count % W == 0) ? W : 1, units = pixel_count / unit;
            const int64_t q = units / N, r = units % N;
            p_d = pixel_begin + unit * (d * q + std::min<int64_t>(d, r)); pc_d = unit * (q + (d < r ? 1 : 0));
        } else {
            const int q = spp / N, r = spp % N;
            n_d = q + (d < r ? 1 : 0); b_d = sample_begin + d * q + std::min(d, r);
        }
        const bool clear_first = d > 0 || fresh;
        if (n_d == 0 || pc_d == 0) {                   // nothing to render here: the buffer still takes part in the reduce
            if (clear_first && cudaMemsetAsync(c->accum, 0, n_floats * 4, c->stream) != cudaSuccess) rcs[d] = set_err(c, PRT_ERR_CUDA, "memset failed");
            return;
        }
        const bool whole = p_d == 0 && pc_d == (int64_t)W * root->cam.height;
        if (clear_first && !whole && cudaMemsetAsync(c->accum, 0, n_floats * 4, c->stream) != cudaSuccess) { rcs[d] = set_err(c, PRT_ERR_CUDA, "memset failed"); return; }
        rcs[d] = wavefront_render(c, n_d, max_depth, seed, b_d, p_d, pc_d, clear_first && whole, /*defer_finish=*/true);
    };
    {
        std::vector<std::thread> th;
        for (int d = 1; d < N; ++d) th.emplace_back(enqueue, d);
        enqueue(0);
        for (auto& t : th) t.join();
    }
    for (int d = 0; d < N; ++d)
        if (rcs[d] != PRT_OK && rc == PRT_OK) rc = d == 0 ? rcs[d] : set_err(root, rcs[d], dev[d]->c.err);
    // ---- the reduce, enqueued behind every device's last wave
    if (rc == PRT_OK) {
        cudaSetDevice(root->device);
        cudaEventRecord(h->evr0, root->stream);
        const bool use_nccl = !h->comms.empty() && h->reduce_mode != 2;
        if (use_nccl) {
            ncclResult_t nr = g_nccl.GroupStart();
            for (int d = 0; d < N && nr == ncclSuccess; ++d) {
                Ctx* c = &dev[d]->c;
                cudaSetDevice(c->device);
                nr = g_nccl.Reduce(c->accum, c->accum, n_floats, ncclFloat, ncclSum, 0, (ncclComm_t)h->comms[d], c->stream);
            }
            ncclResult_t ne = g_nccl.GroupEnd();
            if (nr == ncclSuccess) nr = ne;
            if (nr != ncclSuccess) rc = set_err(root, PRT_ERR_CUDA, std::string("ncclReduce: ") + g_nccl.GetErrorString(nr));
            cudaSetDevice(root->device);
        } else {
            PeerPtrs pp{};
            for (int d = 1; d < N; ++d) {
                Ctx* c = &dev[d]->c;
                cudaSetDevice(c->device);
                cudaEventRecord(h->ev_peer[d - 1], c->stream);
                pp.p[d - 1] = c->accum;
            }
            cudaSetDevice(root->device);
            for (int d = 1; d < N; ++d) cudaStreamWaitEvent(root->stream, h->ev_peer[d - 1], 0);
            const int grid = root->num_sms * 8;
            k_sum_peers<<<grid, 256, 0, root->stream>>>(root->accum, pp, N - 1, n_floats);
        }
        cudaEventRecord(h->evr1, root->stream);
    }
    // ---- wait for every device, gather the counters
    prt_stats tot{};
    for (int d = 0; d < N; ++d) {
        Ctx* c = &dev[d]->c;
        cudaSetDevice(c->device);
        const bool had = c->pend_active;
        int r2 = wavefront_finish(c);
        if (r2 != PRT_OK && rc == PRT_OK) rc = c == root ? r2 : set_err(root, r2, c->err);
        if (!had) continue;
        const prt_stats& s = c->stats;
        tot.rays_total += s.rays_total; tot.paths += s.paths; tot.nodes_visited += s.nodes_visited; tot.prims_tested += s.prims_tested;
        for (int b = 0; b < PRT_MAX_DEPTH; ++b) tot.rays_per_bounce[b] += s.rays_per_bounce[b];
        tot.extend_ms = std::max(tot.extend_ms, s.extend_ms); tot.shade_ms = std::max(tot.shade_ms, s.shade_ms);
        tot.primary_ms = std::max(tot.primary_ms, s.primary_ms); tot.render_ms = std::max(tot.render_ms, s.render_ms);
        tot.kernel_launches += s.kernel_launches; tot.primary_launches += s.primary_launches;
    }
    cudaSetDevice(root->device);
    if (rc == PRT_OK) {
        if (cudaStreamSynchronize(root->stream) != cudaSuccess) rc = set_err(root, PRT_ERR_CUDA, "multi-device reduce failed");
        float red = 0.f, whole_ms = 0.f;
        cudaEventElapsedTime(&red, h->evr0, h->evr1);
        if (root->stats.paths && cudaEventElapsedTime(&whole_ms, root->ev0, h->evr1) == cudaSuccess) tot.render_ms = std::max(tot.render_ms, whole_ms);
        tot.reduce_ms = red;
        tot.kernel_launches += 1;
    }
    tot.build_ms = root->build_ms; tot.n_prims = root->n_prims; tot.n_nodes = root->n_prims > 1 ? root->n_prims - 1 : 0;
    root->stats = tot;
    const bool whole_frame = pixel_begin == 0 && pixel_count == (int64_t)W * root->cam.height;
    root->accum_spp = whole_frame ? base_spp + spp : base_spp;
    return rc;
}

extern "C" {

const char* prt_version(void) { return "parallelraytracing_b200 0.1 (sm_100a)"; }

const char* prt_last_error(const prt_ctx* ctx) { return ctx ? ctx->c.err.c_str() : g_err.c_str(); }

int prt_create(prt_ctx** out, int device) {
    if (!out) return set_err(nullptr, PRT_ERR_INVALID, "prt_create: out is NULL");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return set_err(nullptr, PRT_ERR_NO_DEVICE, std::string("no CUDA device: ") + cudaGetErrorString(e));
    }
    if (device < 0 || device >= ndev) return set_err(nullptr, PRT_ERR_INVALID, "prt_create: bad device ordinal");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return set_err(nullptr, PRT_ERR_CUDA, "cudaGetDeviceProperties failed");
    if (prop.major < 10) {
        char buf[128];
        snprintf(buf, sizeof buf, "device %d is sm_%d%d; this library ships sm_100a code only", device, prop.major, prop.minor);
        return set_err(nullptr, PRT_ERR_NO_DEVICE, buf);
    }
    prt_ctx* h = new (std::nothrow) prt_ctx();
    if (!h) return set_err(nullptr, PRT_ERR_OOM, "host allocation failed");
    Ctx* c = &h->c;
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess) {
        delete h;
        return set_err(nullptr, PRT_ERR_CUDA, "stream/event creation failed");
    }
    *out = h;
    return PRT_OK;
}

int prt_create_multi(prt_ctx** out, const int* device_ids, int n_devices) {
    if (!out) return set_err(nullptr, PRT_ERR_INVALID, "prt_create_multi: out is NULL");
    *out = nullptr;
    if (!device_ids || n_devices < 1 || n_devices > 16) return set_err(nullptr, PRT_ERR_INVALID, "prt_create_multi: need 1..16 device ids");
    for (int i = 0; i < n_devices; ++i)
        for (int j = 0; j < i; ++j)
            if (device_ids[i] == device_ids[j]) return set_err(nullptr, PRT_ERR_INVALID, "prt_create_multi: duplicate device id");
    prt_ctx* root = nullptr;
    int rc = prt_create(&root, device_ids[0]);
    if (rc != PRT_OK) return rc;
    for (int i = 1; i < n_devices; ++i) {
        prt_ctx* p = nullptr;
        rc = prt_create(&p, device_ids[i]);
        if (rc != PRT_OK) { prt_destroy(root); return rc; }
        root->peers.push_back(p);
    }
    if (n_devices > 1) {
        cudaSetDevice(device_ids[0]);
        cudaEventCreate(&root->evr0); cudaEventCreate(&root->evr1);
        root->ev_peer.resize(n_devices - 1);
        for (int i = 1; i < n_devices; ++i) {
            cudaSetDevice(device_ids[i]);
            cudaEventCreateWithFlags(&root->ev_peer[i - 1], cudaEventDisableTiming);
            int can = 0;                                   // peer mappings for the peer-to-peer reduce (NVLink on one box)
            cudaSetDevice(device_ids[0]);
            if (cudaDeviceCanAccessPeer(&can, device_ids[0], device_ids[i]) == cudaSuccess && can) {
                cudaError_t e = cudaDeviceEnablePeerAccess(device_ids[i], 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) can = 0;
                cudaGetLastError();
            }
            if (i == 1) root->p2p_enabled = can != 0; else root->p2p_enabled = root->p2p_enabled && can;
        }
        if (g_nccl.load()) {                               // ncclCommInitAll: one communicator per device of this process
            std::vector<ncclComm_t> comms(n_devices);
            ncclResult_t nr = g_nccl.CommInitAll(comms.data(), n_devices, device_ids);
            if (nr == ncclSuccess) for (ncclComm_t cm : comms) root->comms.push_back((void*)cm);
            else cudaGetLastError();
        }
        if (root->comms.empty() && !root->p2p_enabled) {
            prt_destroy(root);
            return set_err(nullptr, PRT_ERR_UNSUPPORTED, "prt_create_multi: neither libnccl.so.2 nor peer access between the devices is available");
        }
        cudaSetDevice(device_ids[0]);
    }
    *out = root;
    return PRT_OK;
}

void prt_destroy(prt_ctx* ctx) {
    if (!ctx) return;
    for (void* cm : ctx->comms) if (cm) g_nccl.CommDestroy((ncclComm_t)cm);
    ctx->comms.clear();
    for (prt_ctx* p : ctx->peers) prt_destroy(p);
    ctx->peers.clear();
    Ctx* c = &ctx->c;
    cudaSetDevice(c->device);
    for (cudaEvent_t e : ctx->ev_peer) if (e) cudaEventDestroy(e);
    if (ctx->evr0) cudaEventDestroy(ctx->evr0);
    if (ctx->evr1) cudaEventDestroy(ctx->evr1);
    if (c->stream) cudaStreamSynchronize(c->stream);
    DevBuf* bufs[] = {&c->spheres, &c->quads, &c->quad_mat, &c->mats, &c->tri_in_v, &c->tri_in_n, &c->tri_in_mat,
                      &c->prim_id_of_q, &c->aabb_min, &c->leaf_box, &c->scene_bounds, &c->keys[0], &c->keys[1],
                      &c->vals[0], &c->vals[1], &c->block_hist, &c->children, &c->parent, &c->node_min, &c->node_max,
                      &c->flags, &c->nodes, &c->wnodes, &c->wnodes_c, &c->wc_queue, &c->tri_v, &c->tri_n, &c->morton, &c->accum_own, &c->display,
                      &c->q_o[0], &c->q_o[1], &c->q_d[0], &c->q_d[1], &c->q_b[0], &c->q_b[1], &c->hits, &c->contrib,
                      &c->q2_o[0], &c->q2_o[1], &c->q2_d[0], &c->q2_d[1], &c->q2_b[0], &c->q2_b[1], &c->hits2, &c->contrib2,
                      &c->qcount2, &c->qcount, &c->stats_dev, &c->ray_keys[0][0], &c->ray_keys[0][1], &c->ray_keys[1][0],
                      &c->ray_keys[1][1], &c->ray_vals[0][0], &c->ray_vals[0][1], &c->ray_vals[1][0], &c->ray_vals[1][1],
                      &c->ray_hist[0], &c->ray_hist[1], &c->ploc_code, &c->ploc_box, &c->ploc_status, &c->ploc_ctl, &c->raw_v, &c->raw_n, &c->raw_m, &c->tri_bounds, &c->dyn_dev};
    for (DevBuf* b : bufs) dev_free(*b);
    wavefront_drop_graphs(c);
    if (c->dyn_host) cudaFreeHost(c->dyn_host);
    if (c->stats_host) cudaFreeHost(c->stats_host);
    for (int k = 0; k < 2; ++k) { if (c->stage[k]) cudaFreeHost(c->stage[k]); if (c->stage_ev[k]) cudaEventDestroy(c->stage_ev[k]); }
    for (cudaEvent_t e : c->ev_pool) cudaEventDestroy(e);
    for (cudaEvent_t e : {c->ev_res[0], c->ev_res[1], c->ev_join}) if (e) cudaEventDestroy(e);
    if (c->stream2) { cudaStreamSynchronize(c->stream2); cudaStreamDestroy(c->stream2); }
    if (c->up_stream) { cudaStreamSynchronize(c->up_stream); cudaStreamDestroy(c->up_stream); }
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete ctx;
}

// prt_set_scene (build = false) and prt_set_scene_build (build = true: the BVH build runs on the positions while the vertex
// normals and material indices are still crossing PCIe; see prt.h)
static int set_scene_impl(prt_ctx* ctx, const prt_sphere* spheres, int n_spheres, const prt_quad* quads, int n_quads,
                          const float* tri_verts, const float* tri_normals, const int32_t* tri_material,
                          int tri_material_default, int n_tris, int tri_prim_base,
                          const prt_material* materials, int n_materials, const float sky_rgb[3], bool build) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_set_scene: ctx is NULL");
    Ctx* c = &ctx->c;
    if (n_spheres < 0 || n_quads < 0 || n_tris < 0 || n_materials < 0 ||
        (n_spheres && !spheres) || (n_quads && !quads) || (n_tris && !tri_verts) || (n_materials && !materials))
        return set_err(c, PRT_ERR_INVALID, "prt_set_scene: negative count or NULL array");
    if ((int64_t)n_spheres + n_quads + n_tris >= ((int64_t)1 << 28))
        return set_err(c, PRT_ERR_UNSUPPORTED, "prt_set_scene: more than 2^28 primitives");
    PRT_CUDA(c, cudaSetDevice(c->device));
    c->scene_set = false; c->bvh_built = false; c->hint_rays[0] = 0;
    wavefront_drop_graphs(c);
    int rc;
    // spheres
    std::vector<SphereRec> hs((size_t)n_spheres);
    for (int i = 0; i < n_spheres; ++i) {
        const prt_sphere& s = spheres[i];
        hs[i].a = make_float4(s.cx, s.cy, s.cz, s.radius);
        float mat, pid; memcpy(&mat, &s.material, 4); memcpy(&pid, &s.prim_id, 4);
        hs[i].b = make_float4(s.tmin, mat, pid, 0.f);
    }
    if ((rc = upload(c, c->spheres, hs.data(), hs.size() * sizeof(SphereRec)))) return rc;
    // quads
    std::vector<QuadRec> hq((size_t)n_quads);
    std::vector<int> hqm((size_t)n_quads);
    for (int i = 0; i < n_quads; ++i) {
        const prt_quad& q = quads[i];
        float pid; memcpy(&pid, &q.prim_id, 4);
        hq[i].a = make_float4(q.cx, q.cy, q.cz, q.half_w);
        hq[i].b = make_float4(q.ux, q.uy, q.uz, q.half_h);
        hq[i].c = make_float4(q.vx, q.vy, q.vz, q.tmin);
        hq[i].d = make_float4(q.nx, q.ny, q.nz, pid);
        hqm[i] = q.material;
    }
    if ((rc = upload(c, c->quads, hq.data(), hq.size() * sizeof(QuadRec)))) return rc;
    if ((rc = upload(c, c->quad_mat, hqm.data(), hqm.size() * sizeof(int)))) return rc;
    // materials
    std::vector<MatRec> hm((size_t)n_materials);
    for (int i = 0; i < n_materials; ++i) {
        const prt_material& m = materials[i];
        float ty; memcpy(&ty, &m.type, 4);
        hm[i].a = make_float4(m.r, m.g, m.b, m.param);
        hm[i].b = make_float4(ty, 0.f, 0.f, 0.f);
    }
    if ((rc = upload(c, c->mats, hm.data(), hm.size() * sizeof(MatRec)))) return rc;
    // triangles: raw arrays straight to the device (staging buffers live in the context: no cudaMalloc/cudaFree per call),
    // expanded there (k_expand_tris)
    const bool overlap = build && n_tris > 0 && (tri_normals || tri_material);
    if (n_tris > 0) {
        const size_t nt = (size_t)n_tris;
        rc = upload_large(c, c->raw_v, tri_verts, nt * 36);
        if (!rc && tri_normals) rc = overlap ? dev_reserve(c, c->raw_n, nt * 36) : upload_large(c, c->raw_n, tri_normals, nt * 36);
        if (!rc && tri_material) rc = overlap ? dev_reserve(c, c->raw_m, nt * 4) : upload_large(c, c->raw_m, tri_material, nt * 4);
        if (rc) return rc;
    }
    // records the build gathers + every triangle's box and their share of the bounds (k_ingest_tris)
    if ((rc = lbvh_ingest_tris(c, n_spheres, n_quads, n_tris, tri_normals != nullptr, tri_material != nullptr, tri_material_default, overlap))) return rc;
    c->tri_deferred = overlap; c->def_normals = tri_normals != nullptr; c->def_mat = tri_material != nullptr; c->def_mat_default = tri_material_default;
    c->n_spheres = n_spheres; c->n_quads = n_quads; c->n_tris = n_tris; c->n_mats = n_materials;
    c->n_prims = n_spheres + n_quads + n_tris;
    c->tri_prim_base = tri_prim_base;
    if (sky_rgb) { c->sky[0] = sky_rgb[0]; c->sky[1] = sky_rgb[1]; c->sky[2] = sky_rgb[2]; }
    if (!build) {
        PRT_CUDA(c, cudaStreamSynchronize(c->stream));             // the caller's arrays (and the records above) are free again
        PRT_CUDA(c, cudaGetLastError());
        c->scene_set = true;
    } else {
        // The build needs positions only (boxes, Morton codes, sort, hierarchy, wide nodes, vertex gather): it runs on the
        // render stream while a second host thread sends the normals and materials over a stream of its own (a pageable source keeps
        // that thread busy filling the staging ring); k_finish_tris then drops them into their leaf slots.
        c->scene_set = true;
        int up_rc = PRT_OK; cudaError_t up_err = cudaSuccess;
        std::thread uploader;
        if (overlap) {
            if (!c->up_stream) PRT_CUDA(c, cudaStreamCreateWithFlags(&c->up_stream, cudaStreamNonBlocking));
            PRT_CUDA(c, cudaStreamSynchronize(c->stream));         // the staging ring is free (positions are on the device), boxes are done
            uploader = std::thread([&] {
                up_err = cudaSetDevice(c->device);
                if (up_err != cudaSuccess) return;
                const size_t nt = (size_t)n_tris;
                if (tri_normals) up_rc = upload_large(c, c->raw_n, tri_normals, nt * 36, c->up_stream, false);
                if (!up_rc && tri_material) up_rc = upload_large(c, c->raw_m, tri_material, nt * 4, c->up_stream, false);
                if (!up_rc) up_err = cudaStreamSynchronize(c->up_stream);
            });
        }
        c->finish_later = overlap;                              // the normals are still in flight: k_finish_tris runs below, after the join
        rc = lbvh_build(c);
        c->finish_later = false;
        if (uploader.joinable()) uploader.join();
        if (rc == PRT_OK && up_rc != PRT_OK) rc = up_rc;
        if (rc == PRT_OK && up_err != cudaSuccess) rc = set_err(c, PRT_ERR_CUDA, std::string("prt_set_scene_build: upload: ") + cudaGetErrorString(up_err));
        if (rc == PRT_OK && overlap) {
            rc = lbvh_finish_tris(c);
            if (rc == PRT_OK) {
                cudaError_t e = cudaStreamSynchronize(c->stream);
                if (e == cudaSuccess) e = cudaGetLastError();
                if (e != cudaSuccess) rc = set_err(c, PRT_ERR_CUDA, std::string("prt_set_scene_build: ") + cudaGetErrorString(e));
            }
        }
        if (rc != PRT_OK) { c->scene_set = false; c->bvh_built = false; return rc; }
    }
    for (prt_ctx* p : ctx->peers) {                        // multi-device context: the scene is replicated (every device builds its own BVH: identical bits)
        int rp = set_scene_impl(p, spheres, n_spheres, quads, n_quads, tri_verts, tri_normals, tri_material, tri_material_default,
                                n_tris, tri_prim_base, materials, n_materials, sky_rgb, build);
        if (rp != PRT_OK) return set_err(c, rp, p->c.err);
    }
    if (!ctx->peers.empty()) PRT_CUDA(c, cudaSetDevice(c->device));
    return PRT_OK;
}

int prt_set_scene(prt_ctx* ctx, const prt_sphere* spheres, int n_spheres, const prt_quad* quads, int n_quads,
                  const float* tri_verts, const float* tri_normals, const int32_t* tri_material,
                  int tri_material_default, int n_tris, int tri_prim_base,
                  const prt_material* materials, int n_materials, const float sky_rgb[3]) {
    return set_scene_impl(ctx, spheres, n_spheres, quads, n_quads, tri_verts, tri_normals, tri_material, tri_material_default,
                          n_tris, tri_prim_base, materials, n_materials, sky_rgb, false);
}

int prt_set_scene_build(prt_ctx* ctx, const prt_sphere* spheres, int n_spheres, const prt_quad* quads, int n_quads,
                        const float* tri_verts, const float* tri_normals, const int32_t* tri_material,
                        int tri_material_default, int n_tris, int tri_prim_base,
                        const prt_material* materials, int n_materials, const float sky_rgb[3]) {
    return set_scene_impl(ctx, spheres, n_spheres, quads, n_quads, tri_verts, tri_normals, tri_material, tri_material_default,
                          n_tris, tri_prim_base, materials, n_materials, sky_rgb, true);
}

int prt_build_bvh(prt_ctx* ctx) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_build_bvh: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!c->scene_set) return set_err(c, PRT_ERR_INVALID, "prt_build_bvh: call prt_set_scene first");
    PRT_CUDA(c, cudaSetDevice(c->device));
    wavefront_drop_graphs(c);
    int rc = lbvh_build(c);
    for (prt_ctx* p : ctx->peers) {                        // every device builds its own BVH (deterministic: identical bits, no broadcast)
        if (rc != PRT_OK) break;
        rc = prt_build_bvh(p);
        if (rc != PRT_OK) set_err(c, rc, p->c.err);
    }
    if (!ctx->peers.empty()) cudaSetDevice(c->device);
    return rc;
}

int prt_set_camera(prt_ctx* ctx, const prt_camera* cam) {
    if (!ctx || !cam) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_set_camera: NULL argument");
    Ctx* c = &ctx->c;
    if (cam->width <= 0 || cam->height <= 0 || (int64_t)cam->width * cam->height > ((int64_t)1 << 30))
        return set_err(c, PRT_ERR_INVALID, "prt_set_camera: bad resolution");
    PRT_CUDA(c, cudaSetDevice(c->device));
    bool resized = !c->cam_set || cam->width != c->cam.width || cam->height != c->cam.height;
    if (resized && c->accum_external)                      // the caller's buffer has the old size: it must re-bind, not lose samples silently
        return set_err(c, PRT_ERR_INVALID, "prt_set_camera: resolution changed while an external accumulation buffer is bound "
                                           "(call prt_set_accum_buffer(ctx, NULL) or bind a buffer of the new size first)");
    for (prt_ctx* p : ctx->peers) {
        int rp = prt_set_camera(p, cam);
        if (rp != PRT_OK) return set_err(c, rp, p->c.err);
    }
    PRT_CUDA(c, cudaSetDevice(c->device));
    wavefront_drop_graphs(c);
    c->cam = *cam;
    c->cam_set = true;
    if (resized) {
        size_t bytes = (size_t)cam->width * cam->height * 3 * sizeof(float);
        int rc = dev_reserve(c, c->accum_own, bytes);
        if (rc) return rc;
        c->accum = (float*)c->accum_own.p;
        PRT_CUDA(c, cudaMemsetAsync(c->accum, 0, bytes, c->stream));
        PRT_CUDA(c, cudaStreamSynchronize(c->stream));
        c->accum_spp = 0;
    }
    return PRT_OK;
}

int prt_render(prt_ctx* ctx, int spp, int max_depth, uint64_t seed, int sample_begin) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_render: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!argcheck_render(c, "prt_render", spp, max_depth, sample_begin, 0)) return PRT_ERR_INVALID;
    if (spp == 0) return PRT_OK;
    PRT_CUDA(c, cudaSetDevice(c->device));
    if (max_depth == 0) { c->accum_spp += spp; return PRT_OK; }   // TraceRay(depth<=0) = 0, renderer.cpp:61
    const int64_t P = (int64_t)c->cam.width * c->cam.height;
    if (!ctx->peers.empty()) return multi_render(ctx, spp, max_depth, seed, sample_begin, 0, P, false);
    return wavefront_render(c, spp, max_depth, seed, sample_begin, 0, P);
}

int prt_render_fresh(prt_ctx* ctx, int spp, int max_depth, uint64_t seed, int sample_begin) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_render_fresh: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!argcheck_render(c, "prt_render_fresh", spp, max_depth, sample_begin, 1)) return PRT_ERR_INVALID;
    PRT_CUDA(c, cudaSetDevice(c->device));
    if (max_depth == 0) {                                          // TraceRay(depth<=0) = 0, renderer.cpp:61
        PRT_CUDA(c, cudaMemsetAsync(c->accum, 0, (size_t)c->cam.width * c->cam.height * 12, c->stream));
        PRT_CUDA(c, cudaStreamSynchronize(c->stream));
        c->accum_spp = spp;
        return PRT_OK;
    }
    const int64_t P = (int64_t)c->cam.width * c->cam.height;
    if (!ctx->peers.empty()) return multi_render(ctx, spp, max_depth, seed, sample_begin, 0, P, true);
    return wavefront_render(c, spp, max_depth, seed, sample_begin, 0, P, true);
}

int prt_render_region(prt_ctx* ctx, int spp, int max_depth, uint64_t seed, int sample_begin, int pixel_begin, int pixel_count) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_render_region: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!argcheck_render(c, "prt_render_region", spp, max_depth, sample_begin, 0)) return PRT_ERR_INVALID;
    const int64_t P = (int64_t)c->cam.width * c->cam.height;
    if (pixel_begin < 0 || pixel_count < 0 || (int64_t)pixel_begin + pixel_count > P)
        return set_err(c, PRT_ERR_INVALID, "prt_render_region: bad spp / max_depth / sample_begin / pixel range");
    if (spp == 0 || pixel_count == 0) return PRT_OK;
    PRT_CUDA(c, cudaSetDevice(c->device));
    const bool whole = pixel_begin == 0 && pixel_count == P;
    if (max_depth == 0) { if (whole) c->accum_spp += spp; return PRT_OK; }
    if (!ctx->peers.empty()) return multi_render(ctx, spp, max_depth, seed, sample_begin, pixel_begin, pixel_count, false);
    const int before = c->accum_spp;
    int rc = wavefront_render(c, spp, max_depth, seed, sample_begin, pixel_begin, pixel_count);
    if (!whole) c->accum_spp = before;
    return rc;
}

int prt_wait_event(prt_ctx* ctx, void* cuda_event) {
    if (!ctx || !cuda_event) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_wait_event: NULL argument");
    Ctx* c = &ctx->c;
    PRT_CUDA(c, cudaSetDevice(c->device));
    PRT_CUDA(c, cudaStreamWaitEvent(c->stream, (cudaEvent_t)cuda_event, 0));
    return PRT_OK;
}

int prt_clear_accum(prt_ctx* ctx) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_clear_accum: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!c->cam_set) return set_err(c, PRT_ERR_INVALID, "prt_clear_accum: no camera/film yet");
    PRT_CUDA(c, cudaSetDevice(c->device));
    PRT_CUDA(c, cudaMemsetAsync(c->accum, 0, (size_t)c->cam.width * c->cam.height * 12, c->stream));
    PRT_CUDA(c, cudaStreamSynchronize(c->stream));
    c->accum_spp = 0;
    return PRT_OK;
}

int prt_scale_accum(prt_ctx* ctx, float s) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_scale_accum: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!c->cam_set) return set_err(c, PRT_ERR_INVALID, "prt_scale_accum: no film yet");
    PRT_CUDA(c, cudaSetDevice(c->device));
    return film_scale_accum(c, s);
}

float* prt_accum_device_ptr(prt_ctx* ctx) { return ctx ? ctx->c.accum : nullptr; }

int prt_set_accum_buffer(prt_ctx* ctx, float* device_rgb) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_set_accum_buffer: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!c->cam_set) return set_err(c, PRT_ERR_INVALID, "prt_set_accum_buffer: call prt_set_camera first");
    cudaSetDevice(c->device); wavefront_drop_graphs(c);
    if (device_rgb) { c->accum = device_rgb; c->accum_external = true; }
    else { c->accum = (float*)c->accum_own.p; c->accum_external = false; }
    c->accum_spp = 0;
    return PRT_OK;
}

int prt_read_accum(prt_ctx* ctx, float* host_rgb) {
    if (!ctx || !host_rgb) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_read_accum: NULL argument");
    Ctx* c = &ctx->c;
    if (!c->cam_set) return set_err(c, PRT_ERR_INVALID, "prt_read_accum: no film yet");
    PRT_CUDA(c, cudaSetDevice(c->device));
    PRT_CUDA(c, cudaMemcpyAsync(host_rgb, c->accum, (size_t)c->cam.width * c->cam.height * 12, cudaMemcpyDeviceToHost, c->stream));
    PRT_CUDA(c, cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

int prt_accum_spp(const prt_ctx* ctx) { return ctx ? ctx->c.accum_spp : 0; }

int prt_set_accum_spp(prt_ctx* ctx, int spp) {
    if (!ctx || spp < 0) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_set_accum_spp: bad argument");
    ctx->c.accum_spp = spp;
    return PRT_OK;
}

int prt_resolve_display(prt_ctx* ctx, float exposure, float gamma, uint8_t* host_rgba) {
    if (!ctx || !host_rgba) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_resolve_display: NULL argument");
    Ctx* c = &ctx->c;
    if (!c->cam_set) return set_err(c, PRT_ERR_INVALID, "prt_resolve_display: no film yet");
    PRT_CUDA(c, cudaSetDevice(c->device));
    return film_resolve_display(c, exposure, gamma, host_rgba);
}

int prt_trace_batch(prt_ctx* ctx, const prt_ray* rays, int n, prt_hit* out, int use_bvh) {
    if (!ctx || (n > 0 && (!rays || !out))) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_trace_batch: NULL argument");
    Ctx* c = &ctx->c;
    if (n < 0 || n > (1 << 26)) return set_err(c, PRT_ERR_INVALID, "prt_trace_batch: bad n");
    if (!c->scene_set || !c->bvh_built) return set_err(c, PRT_ERR_INVALID, "prt_trace_batch: scene/BVH not ready");
    PRT_CUDA(c, cudaSetDevice(c->device));
    return wavefront_trace_batch(c, rays, n, out, use_bvh);
}

int prt_debug_scatter(prt_ctx* ctx, const prt_material* mats, const float* in_dirs, const float* normals, const int32_t* fronts,
                      const uint32_t* keys, uint64_t seed, int n, float* out) {
    if (!ctx || n < 0 || (n > 0 && (!mats || !in_dirs || !normals || !fronts || !keys || !out)))
        return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_debug_scatter: NULL argument or negative count");
    PRT_CUDA(&ctx->c, cudaSetDevice(ctx->c.device));
    return wavefront_debug_scatter(&ctx->c, mats, in_dirs, normals, fronts, keys, seed, n, out);
}

int prt_debug_check_flags(prt_ctx* ctx) {
    if (!ctx) return PRT_ERR_INVALID;
#if PRT_CHECKED
    int flags = 0;
    for (prt_ctx* d : all_devices(ctx)) {
        cudaSetDevice(d->c.device);
        cudaDeviceSynchronize();
        flags |= (int)(wavefront_check_flags() | lbvh_check_flags());
    }
    cudaSetDevice(ctx->c.device);
    return flags | 0x40000000;                 // bit 30: this IS the checked library
#else
    return 0;
#endif
}

void* prt_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 16, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void prt_host_free(void* p) { if (p) cudaFreeHost(p); }

int prt_num_prims(const prt_ctx* ctx) { return ctx ? ctx->c.n_prims : 0; }

int prt_debug_lbvh(prt_ctx* ctx, uint32_t* morton, uint32_t* sorted_prim, int32_t* node_children,
                   int32_t* node_parent, float* node_bounds, float* scene_bounds) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_debug_lbvh: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!c->bvh_built) return set_err(c, PRT_ERR_INVALID, "prt_debug_lbvh: BVH not built");
    PRT_CUDA(c, cudaSetDevice(c->device));
    const int n = c->n_prims;
    if (scene_bounds) {
        int ob[6];
        PRT_CUDA(c, cudaMemcpy(ob, c->scene_bounds.p, sizeof ob, cudaMemcpyDeviceToHost));
        for (int k = 0; k < 6; ++k) { int i = ob[k] >= 0 ? ob[k] : ob[k] ^ 0x7fffffff; memcpy(&scene_bounds[k], &i, 4); }
    }
    if (n == 0) return PRT_OK;
    if (morton) PRT_CUDA(c, cudaMemcpy(morton, c->morton.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    if (sorted_prim) PRT_CUDA(c, cudaMemcpy(sorted_prim, c->vals[c->sorted_slot].p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    if (n > 1) {
        if (node_children) PRT_CUDA(c, cudaMemcpy(node_children, c->children.p, (size_t)(n - 1) * 8, cudaMemcpyDeviceToHost));
        if (node_parent) PRT_CUDA(c, cudaMemcpy(node_parent, c->parent.p, (size_t)(n - 1) * 4, cudaMemcpyDeviceToHost));
        if (node_bounds) {
            std::vector<BvhNode> h((size_t)n - 1);
            PRT_CUDA(c, cudaMemcpy(h.data(), c->nodes.p, h.size() * sizeof(BvhNode), cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < h.size(); ++i) {
                float* o = node_bounds + 12 * i;
                o[0] = h[i].lmin.x; o[1] = h[i].lmin.y; o[2] = h[i].lmin.z; o[3] = h[i].lmax.x; o[4] = h[i].lmax.y; o[5] = h[i].lmax.z;
                o[6] = h[i].rmin.x; o[7] = h[i].rmin.y; o[8] = h[i].rmin.z; o[9] = h[i].rmax.x; o[10] = h[i].rmax.y; o[11] = h[i].rmax.z;
            }
        }
    }
    return PRT_OK;
}

int prt_debug_ploc_trace(prt_ctx* ctx, int32_t* counts, uint32_t* ns, int max_rounds) {
    if (!ctx || !counts || !ns) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_debug_ploc_trace: NULL argument");
    const Ctx* c = &ctx->c;
    const int n = std::min(max_rounds, c->ploc_rounds + 1);
    for (int r = 0; r < n; ++r) { counts[r] = c->ploc_trace[16 + 2 * r]; ns[r] = (uint32_t)c->ploc_trace[16 + 2 * r + 1]; }
    return n;
}

int prt_debug_wide(prt_ctx* ctx, uint32_t* words) {
    if (!ctx || !words) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_debug_wide: NULL argument");
    Ctx* c = &ctx->c;
    if (!c->bvh_built || !c->wide_built) return set_err(c, PRT_ERR_INVALID, "prt_debug_wide: wide nodes not built");
    PRT_CUDA(c, cudaSetDevice(c->device));
    PRT_CUDA(c, cudaMemcpy(words, c->wnodes.p, (size_t)(c->n_prims - 1) * sizeof(WideNode), cudaMemcpyDeviceToHost));
    return PRT_OK;
}

int prt_debug_wide_compact(prt_ctx* ctx, uint32_t* words, int capacity_nodes, int* n_nodes) {
    if (!ctx || !n_nodes) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_debug_wide_compact: NULL argument");
    Ctx* c = &ctx->c;
    if (!c->bvh_built || !c->wide_built || !c->wide_compacted) return set_err(c, PRT_ERR_INVALID, "prt_debug_wide_compact: no compact wide nodes (option wide_compact, scenes with wide nodes)");
    *n_nodes = c->wide_live;
    if (!words) return PRT_OK;
    if (capacity_nodes < c->wide_live) return set_err(c, PRT_ERR_INVALID, "prt_debug_wide_compact: buffer too small");
    PRT_CUDA(c, cudaSetDevice(c->device));
    PRT_CUDA(c, cudaMemcpy(words, c->wnodes_c.p, (size_t)c->wide_live * sizeof(WideNode), cudaMemcpyDeviceToHost));
    return PRT_OK;
}

int prt_enable_counters(prt_ctx* ctx, int on) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_enable_counters: ctx is NULL");
    ctx->c.counters_on = on ? 1 : 0;
    for (prt_ctx* p : ctx->peers) p->c.counters_on = on ? 1 : 0;
    // (the graph key holds the counter switch: no invalidation needed)
    return PRT_OK;
}

int prt_get_stats(prt_ctx* ctx, prt_stats* out) {
    if (!ctx || !out) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_get_stats: NULL argument");
    ctx->c.stats.build_ms = ctx->c.build_ms;
    ctx->c.stats.n_prims = ctx->c.n_prims;
    ctx->c.stats.n_nodes = ctx->c.n_prims > 1 ? ctx->c.n_prims - 1 : 0;
    ctx->c.stats.n_devices = 1 + (int)ctx->peers.size();
    ctx->c.stats.bvh_builder = ctx->c.builder_used; ctx->c.stats.bvh_height = ctx->c.bvh_height; ctx->c.stats.ploc_rounds = ctx->c.ploc_rounds;
    *out = ctx->c.stats;
    return PRT_OK;
}

int prt_set_option(prt_ctx* ctx, const char* key, int64_t value) {
    if (!ctx || !key) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_set_option: NULL argument");
    Ctx* c = &ctx->c;
    if (!strcmp(key, "multi_partition")) {
        if (value != PRT_PARTITION_SAMPLES && value != PRT_PARTITION_TILES) return set_err(c, PRT_ERR_INVALID, "multi_partition: 0 (samples) or 1 (tiles)");
        ctx->partition = (int)value; return PRT_OK;
    }
    if (!strcmp(key, "multi_reduce")) {
        if (value < 0 || value > 2) return set_err(c, PRT_ERR_INVALID, "multi_reduce: 0 auto, 1 NCCL, 2 peer-to-peer kernel");
        if (value == 1 && !ctx->peers.empty() && ctx->comms.empty()) return set_err(c, PRT_ERR_UNSUPPORTED, "multi_reduce=1: libnccl.so.2 could not be loaded");
        if (value == 2 && !ctx->peers.empty() && !ctx->p2p_enabled) return set_err(c, PRT_ERR_UNSUPPORTED, "multi_reduce=2: no peer access between the devices");
        ctx->reduce_mode = (int)value; return PRT_OK;
    }
    for (prt_ctx* p : ctx->peers) {                        // every other option is per device: replicate
        int rp = prt_set_option(p, key, value);
        if (rp != PRT_OK) return set_err(c, rp, p->c.err);
    }
    cudaSetDevice(c->device); wavefront_drop_graphs(c);               // any option may change what a call enqueues
    if (!strcmp(key, "use_graph")) { c->use_graph = value ? 1 : 0; return PRT_OK; }
    if (!strcmp(key, "fast_math")) { c->fast_math = value ? 1 : 0; return PRT_OK; }
    if (!strcmp(key, "wave_paths")) { if (value < 0) return set_err(c, PRT_ERR_INVALID, "wave_paths < 0"); c->wave_paths_opt = value; }
    else if (!strcmp(key, "ctas_per_sm")) { if (value < 1 || value > 32) return set_err(c, PRT_ERR_INVALID, "ctas_per_sm out of range"); c->ctas_per_sm = (int)value; }
    else if (!strcmp(key, "tail_bounce")) c->tail_bounce = value < 0 ? 0 : value;
    else if (!strcmp(key, "tail_max_paths")) c->tail_max_paths = value < 0 ? 0 : value;
    else if (!strcmp(key, "tail_adaptive")) c->tail_adaptive = value ? 1 : 0;
    else if (!strcmp(key, "tail_pool")) c->tail_pool = value ? 1 : 0;
    else if (!strcmp(key, "tail_rays")) c->tail_rays = value < 0 ? 0 : value;
    else if (!strcmp(key, "morton_centroid")) c->morton_centroid = value ? 1 : 0;
    else if (!strcmp(key, "sort_rays")) c->sort_rays = value ? 1 : 0;
    else if (!strcmp(key, "sort_key_bits")) c->sort_key_bits = (int)value;
    else if (!strcmp(key, "sort_min_rays")) c->sort_min_rays = value < 0 ? 0 : (int)value;
    else if (!strcmp(key, "tail_rays_staged")) c->tail_rays_staged = value < 0 ? 0 : value;
    else if (!strcmp(key, "fused")) c->fused = value < 0 ? -1 : (value ? 1 : 0);
    else if (!strcmp(key, "fused_max_prims")) c->fused_max_prims = (int)value;
    else if (!strcmp(key, "brute_max_prims")) c->brute_max_prims = (int)value;
    else if (!strcmp(key, "bvh_wide")) c->bvh_wide = value < 0 ? -1 : (value ? 1 : 0);   // takes effect at the next prt_build_bvh
    else if (!strcmp(key, "wide_min_prims")) c->wide_min_prims = (int)value;
    else if (!strcmp(key, "wide_compact")) c->wide_compact = value ? 1 : 0;                      // takes effect at the next prt_build_bvh
    else if (!strcmp(key, "bvh_builder")) c->bvh_builder = value ? 1 : 0;                        // takes effect at the next prt_build_bvh
    else if (!strcmp(key, "ploc_radius")) { if (value < 1 || value > 8) return set_err(c, PRT_ERR_INVALID, "ploc_radius out of range (1..8)"); c->ploc_radius = (int)value; }
    else if (!strcmp(key, "overlap")) c->overlap = value ? 1 : 0;
    else if (!strcmp(key, "persistent")) c->persistent = value ? 1 : 0;
    else if (!strcmp(key, "pt_refill")) { if (value < 0 || value > 32) return set_err(c, PRT_ERR_INVALID, "pt_refill out of range"); c->pt_refill = (int)value; }
    else if (!strcmp(key, "pt_refill_wide")) { if (value < 0 || value > 32) return set_err(c, PRT_ERR_INVALID, "pt_refill_wide out of range"); c->pt_refill_wide = (int)value; }
    else if (!strcmp(key, "rr_start")) c->rr_start = value < 0 ? -1 : (int)value;                 // -1 = off (reference estimator)
    else if (!strcmp(key, "firefly_clamp_milli")) c->firefly_clamp = value <= 0 ? 0.f : (float)value * 1e-3f;   // 0 = off
    else if (!strcmp(key, "stage_timing")) c->stage_timing = value ? 1 : 0;
    else if (!strcmp(key, "check_stack_limit")) { cudaSetDevice(c->device); wavefront_set_stack_limit((int)value); }   // checked library only (no-op otherwise)
    else return set_err(c, PRT_ERR_INVALID, std::string("unknown option: ") + key);
    return PRT_OK;
}

}  // extern "C"
