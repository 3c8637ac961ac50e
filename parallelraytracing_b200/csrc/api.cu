// api.cu — the extern "C" boundary declared in include/prt.h.
// Replaces the host half of a reference backend: Init's allocation/upload
// (src/backend/cuda_wavefront/renderer.cu:351-370, soa.cpp:37-114), SetCamera (:534-547)
// and ProgressiveRender's driver loop (:372-433).
#include "prt_internal.h"
#include <cstring>
#include <cstdio>

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

// Triangle ingest on the device: the caller's packed arrays (9 floats per triangle, optional 9 floats of vertex
// normals, optional material index) are copied as they are and expanded here into the 16-byte planes the build
// reads. (Doing this on the host cost 2 s per Init for 10 M triangles: a 960 MB staging vector and a scalar loop.)
// Flat normal when none is given: cross(v1 - v0, v2 - v0), unnormalised (the shading normalises it).
__global__ void k_expand_tris(int n, const float* __restrict__ v, const float* __restrict__ nrm, const int* __restrict__ mat,
                              int mat_default, float4* __restrict__ out_v, float4* __restrict__ out_n) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    float a[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) a[k] = v[9 * (size_t)t + k];
#pragma unroll
    for (int k = 0; k < 3; ++k)                                        // 96-byte record {v0, v1 | material, v2, n0, n1, n2}
        out_v[6 * (size_t)t + k] = make_float4(a[3 * k], a[3 * k + 1], a[3 * k + 2], k == 1 ? __int_as_float(mat ? mat[t] : mat_default) : 0.f);
    if (nrm) {
#pragma unroll
        for (int k = 0; k < 3; ++k)
            out_n[6 * (size_t)t + k] = make_float4(nrm[9 * (size_t)t + 3 * k], nrm[9 * (size_t)t + 3 * k + 1], nrm[9 * (size_t)t + 3 * k + 2], 0.f);
    } else {
        float e1x = a[3] - a[0], e1y = a[4] - a[1], e1z = a[5] - a[2], e2x = a[6] - a[0], e2y = a[7] - a[1], e2z = a[8] - a[2];
        float nx = e1y * e2z - e2y * e1z, ny = e1z * e2x - e2z * e1x, nz = e1x * e2y - e2x * e1y;
#pragma unroll
        for (int k = 0; k < 3; ++k) out_n[6 * (size_t)t + k] = make_float4(nx, ny, nz, 0.f);
    }
}

static int upload(Ctx* c, DevBuf& b, const void* host, size_t bytes) {
    int rc = dev_reserve(c, b, bytes);
    if (rc) return rc;
    if (bytes) PRT_CUDA(c, cudaMemcpyAsync(b.p, host, bytes, cudaMemcpyHostToDevice, c->stream));
    return PRT_OK;
}

}  // namespace prt

using namespace prt;

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

void prt_destroy(prt_ctx* ctx) {
    if (!ctx) return;
    Ctx* c = &ctx->c;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    DevBuf* bufs[] = {&c->spheres, &c->quads, &c->quad_mat, &c->mats, &c->tri_in_v, &c->tri_in_n, &c->tri_in_mat,
                      &c->prim_id_of_q, &c->aabb_min, &c->leaf_box, &c->scene_bounds, &c->keys[0], &c->keys[1],
                      &c->vals[0], &c->vals[1], &c->block_hist, &c->children, &c->parent, &c->node_min, &c->node_max,
                      &c->flags, &c->nodes, &c->wnodes, &c->tri_v, &c->tri_n, &c->morton, &c->accum_own, &c->display,
                      &c->q_o[0], &c->q_o[1], &c->q_d[0], &c->q_d[1], &c->q_b[0], &c->q_b[1], &c->hits, &c->contrib,
                      &c->q2_o[0], &c->q2_o[1], &c->q2_d[0], &c->q2_d[1], &c->q2_b[0], &c->q2_b[1], &c->hits2, &c->contrib2,
                      &c->qcount2, &c->qcount, &c->stats_dev, &c->ray_keys[0][0], &c->ray_keys[0][1], &c->ray_keys[1][0],
                      &c->ray_keys[1][1], &c->ray_vals[0][0], &c->ray_vals[0][1], &c->ray_vals[1][0], &c->ray_vals[1][1],
                      &c->ray_hist[0], &c->ray_hist[1], &c->ploc_code, &c->ploc_box, &c->ploc_status, &c->ploc_ctl};
    for (DevBuf* b : bufs) dev_free(*b);
    if (c->stats_host) cudaFreeHost(c->stats_host);
    for (cudaEvent_t e : c->ev_pool) cudaEventDestroy(e);
    for (cudaEvent_t e : {c->ev_res[0], c->ev_res[1], c->ev_join}) if (e) cudaEventDestroy(e);
    if (c->stream2) { cudaStreamSynchronize(c->stream2); cudaStreamDestroy(c->stream2); }
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete ctx;
}

int prt_set_scene(prt_ctx* ctx, const prt_sphere* spheres, int n_spheres, const prt_quad* quads, int n_quads,
                  const float* tri_verts, const float* tri_normals, const int32_t* tri_material,
                  int tri_material_default, int n_tris, int tri_prim_base,
                  const prt_material* materials, int n_materials, const float sky_rgb[3]) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_set_scene: ctx is NULL");
    Ctx* c = &ctx->c;
    if (n_spheres < 0 || n_quads < 0 || n_tris < 0 || n_materials < 0 ||
        (n_spheres && !spheres) || (n_quads && !quads) || (n_tris && !tri_verts) || (n_materials && !materials))
        return set_err(c, PRT_ERR_INVALID, "prt_set_scene: negative count or NULL array");
    if ((int64_t)n_spheres + n_quads + n_tris >= ((int64_t)1 << 28))
        return set_err(c, PRT_ERR_UNSUPPORTED, "prt_set_scene: more than 2^28 primitives");
    PRT_CUDA(c, cudaSetDevice(c->device));
    c->scene_set = false; c->bvh_built = false; c->hint_rays[0] = 0;
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
    // triangles: raw arrays straight to the device, expanded there (k_expand_tris)
    if (n_tris > 0) {
        const size_t nt = (size_t)n_tris;
        DevBuf raw_v, raw_n, raw_m;
        rc = upload(c, raw_v, tri_verts, nt * 36);
        if (!rc && tri_normals) rc = upload(c, raw_n, tri_normals, nt * 36);
        if (!rc && tri_material) rc = upload(c, raw_m, tri_material, nt * 4);
        if (!rc) rc = dev_reserve(c, c->tri_in_v, nt * 96);
        if (!rc) {
            k_expand_tris<<<(unsigned)((nt + 255) / 256), 256, 0, c->stream>>>(n_tris, (const float*)raw_v.p, (const float*)raw_n.p,
                (const int*)raw_m.p, tri_material_default, (float4*)c->tri_in_v.p, (float4*)c->tri_in_v.p + 3);
        }
        cudaError_t e = cudaStreamSynchronize(c->stream);
        dev_free(raw_v); dev_free(raw_n); dev_free(raw_m);
        if (rc) return rc;
        if (e != cudaSuccess) return set_err(c, PRT_ERR_CUDA, std::string("prt_set_scene: ") + cudaGetErrorString(e));
    }
    PRT_CUDA(c, cudaStreamSynchronize(c->stream));
    c->n_spheres = n_spheres; c->n_quads = n_quads; c->n_tris = n_tris; c->n_mats = n_materials;
    c->n_prims = n_spheres + n_quads + n_tris;
    c->tri_prim_base = tri_prim_base;
    if (sky_rgb) { c->sky[0] = sky_rgb[0]; c->sky[1] = sky_rgb[1]; c->sky[2] = sky_rgb[2]; }
    c->scene_set = true;
    return PRT_OK;
}

int prt_build_bvh(prt_ctx* ctx) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_build_bvh: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!c->scene_set) return set_err(c, PRT_ERR_INVALID, "prt_build_bvh: call prt_set_scene first");
    PRT_CUDA(c, cudaSetDevice(c->device));
    return lbvh_build(c);
}

int prt_set_camera(prt_ctx* ctx, const prt_camera* cam) {
    if (!ctx || !cam) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_set_camera: NULL argument");
    Ctx* c = &ctx->c;
    if (cam->width <= 0 || cam->height <= 0 || (int64_t)cam->width * cam->height > ((int64_t)1 << 30))
        return set_err(c, PRT_ERR_INVALID, "prt_set_camera: bad resolution");
    PRT_CUDA(c, cudaSetDevice(c->device));
    bool resized = !c->cam_set || cam->width != c->cam.width || cam->height != c->cam.height;
    c->cam = *cam;
    c->cam_set = true;
    if (resized) {
        if (c->accum_external) { c->accum_external = false; c->accum = nullptr; }
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
    if (!c->scene_set || !c->bvh_built) return set_err(c, PRT_ERR_INVALID, "prt_render: scene/BVH not ready (prt_set_scene + prt_build_bvh)");
    if (!c->cam_set) return set_err(c, PRT_ERR_INVALID, "prt_render: call prt_set_camera first");
    if (spp < 0 || max_depth < 0 || max_depth > PRT_MAX_DEPTH || sample_begin < 0)
        return set_err(c, PRT_ERR_INVALID, "prt_render: bad spp / max_depth / sample_begin");
    if (spp == 0) return PRT_OK;
    PRT_CUDA(c, cudaSetDevice(c->device));
    if (max_depth == 0) { c->accum_spp += spp; return PRT_OK; }   // TraceRay(depth<=0) = 0, renderer.cpp:61
    return wavefront_render(c, spp, max_depth, seed, sample_begin, 0, (int64_t)c->cam.width * c->cam.height);
}

int prt_render_fresh(prt_ctx* ctx, int spp, int max_depth, uint64_t seed, int sample_begin) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_render_fresh: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!c->scene_set || !c->bvh_built) return set_err(c, PRT_ERR_INVALID, "prt_render_fresh: scene/BVH not ready (prt_set_scene + prt_build_bvh)");
    if (!c->cam_set) return set_err(c, PRT_ERR_INVALID, "prt_render_fresh: call prt_set_camera first");
    if (spp < 1 || max_depth < 0 || max_depth > PRT_MAX_DEPTH || sample_begin < 0)
        return set_err(c, PRT_ERR_INVALID, "prt_render_fresh: bad spp / max_depth / sample_begin");
    PRT_CUDA(c, cudaSetDevice(c->device));
    if (max_depth == 0) {                                          // TraceRay(depth<=0) = 0, renderer.cpp:61
        PRT_CUDA(c, cudaMemsetAsync(c->accum, 0, (size_t)c->cam.width * c->cam.height * 12, c->stream));
        PRT_CUDA(c, cudaStreamSynchronize(c->stream));
        c->accum_spp = spp;
        return PRT_OK;
    }
    return wavefront_render(c, spp, max_depth, seed, sample_begin, 0, (int64_t)c->cam.width * c->cam.height, true);
}

int prt_render_region(prt_ctx* ctx, int spp, int max_depth, uint64_t seed, int sample_begin, int pixel_begin, int pixel_count) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_render_region: ctx is NULL");
    Ctx* c = &ctx->c;
    if (!c->scene_set || !c->bvh_built) return set_err(c, PRT_ERR_INVALID, "prt_render_region: scene/BVH not ready (prt_set_scene + prt_build_bvh)");
    if (!c->cam_set) return set_err(c, PRT_ERR_INVALID, "prt_render_region: call prt_set_camera first");
    const int64_t P = (int64_t)c->cam.width * c->cam.height;
    if (spp < 0 || max_depth < 0 || max_depth > PRT_MAX_DEPTH || sample_begin < 0 || pixel_begin < 0 || pixel_count < 0 ||
        (int64_t)pixel_begin + pixel_count > P)
        return set_err(c, PRT_ERR_INVALID, "prt_render_region: bad spp / max_depth / sample_begin / pixel range");
    if (spp == 0 || pixel_count == 0) return PRT_OK;
    PRT_CUDA(c, cudaSetDevice(c->device));
    const bool whole = pixel_begin == 0 && pixel_count == P;
    if (max_depth == 0) { if (whole) c->accum_spp += spp; return PRT_OK; }
    const int before = c->accum_spp;
    int rc = wavefront_render(c, spp, max_depth, seed, sample_begin, pixel_begin, pixel_count);
    if (!whole) c->accum_spp = before;
    return rc;
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

int prt_debug_wide(prt_ctx* ctx, uint32_t* words) {
    if (!ctx || !words) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_debug_wide: NULL argument");
    Ctx* c = &ctx->c;
    if (!c->bvh_built || !c->wide_built) return set_err(c, PRT_ERR_INVALID, "prt_debug_wide: wide nodes not built");
    PRT_CUDA(c, cudaSetDevice(c->device));
    PRT_CUDA(c, cudaMemcpy(words, c->wnodes.p, (size_t)(c->n_prims - 1) * sizeof(WideNode), cudaMemcpyDeviceToHost));
    return PRT_OK;
}

int prt_enable_counters(prt_ctx* ctx, int on) {
    if (!ctx) return set_err(nullptr, PRT_ERR_INVALID, "prt_enable_counters: ctx is NULL");
    ctx->c.counters_on = on ? 1 : 0;
    return PRT_OK;
}

int prt_get_stats(prt_ctx* ctx, prt_stats* out) {
    if (!ctx || !out) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_get_stats: NULL argument");
    ctx->c.stats.build_ms = ctx->c.build_ms;
    ctx->c.stats.n_prims = ctx->c.n_prims;
    ctx->c.stats.n_nodes = ctx->c.n_prims > 1 ? ctx->c.n_prims - 1 : 0;
    ctx->c.stats.bvh_builder = ctx->c.builder_used; ctx->c.stats.bvh_height = ctx->c.bvh_height; ctx->c.stats.ploc_rounds = ctx->c.ploc_rounds;
    *out = ctx->c.stats;
    return PRT_OK;
}

int prt_set_option(prt_ctx* ctx, const char* key, int64_t value) {
    if (!ctx || !key) return set_err(ctx ? &ctx->c : nullptr, PRT_ERR_INVALID, "prt_set_option: NULL argument");
    Ctx* c = &ctx->c;
    if (!strcmp(key, "wave_paths")) { if (value < 0) return set_err(c, PRT_ERR_INVALID, "wave_paths < 0"); c->wave_paths_opt = value; }
    else if (!strcmp(key, "ctas_per_sm")) { if (value < 1 || value > 32) return set_err(c, PRT_ERR_INVALID, "ctas_per_sm out of range"); c->ctas_per_sm = (int)value; }
    else if (!strcmp(key, "tail_bounce")) c->tail_bounce = value < 0 ? 0 : value;
    else if (!strcmp(key, "tail_max_paths")) c->tail_max_paths = value < 0 ? 0 : value;
    else if (!strcmp(key, "tail_adaptive")) c->tail_adaptive = value ? 1 : 0;
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
    else if (!strcmp(key, "bvh_builder")) c->bvh_builder = value ? 1 : 0;                        // takes effect at the next prt_build_bvh
    else if (!strcmp(key, "ploc_radius")) { if (value < 1 || value > 8) return set_err(c, PRT_ERR_INVALID, "ploc_radius out of range (1..8)"); c->ploc_radius = (int)value; }
    else if (!strcmp(key, "overlap")) c->overlap = value ? 1 : 0;
    else if (!strcmp(key, "persistent")) c->persistent = value ? 1 : 0;
    else if (!strcmp(key, "pt_refill")) { if (value < 0 || value > 32) return set_err(c, PRT_ERR_INVALID, "pt_refill out of range"); c->pt_refill = (int)value; }
    else if (!strcmp(key, "pt_refill_wide")) { if (value < 0 || value > 32) return set_err(c, PRT_ERR_INVALID, "pt_refill_wide out of range"); c->pt_refill_wide = (int)value; }
    else if (!strcmp(key, "rr_start")) c->rr_start = value < 0 ? -1 : (int)value;                 // -1 = off (reference estimator)
    else if (!strcmp(key, "firefly_clamp_milli")) c->firefly_clamp = value <= 0 ? 0.f : (float)value * 1e-3f;   // 0 = off
    else if (!strcmp(key, "stage_timing")) c->stage_timing = value ? 1 : 0;
    else return set_err(c, PRT_ERR_INVALID, std::string("unknown option: ") + key);
    return PRT_OK;
}

}  // extern "C"
