/*
 * ref_harness.cpp — C-ABI harness around the UNMODIFIED reference CPU path.
 * TEST INFRASTRUCTURE ONLY (oracle/): linked with the reference's own sources
 *   src/core/{scene,primitive,renderer,mesh}.cpp, src/core/film.cu, src/backend/cpu/renderer.cpp,
 *   external/tinyply/source/tinyply.cpp
 * compiled where they lie under /root/reference by oracle/Makefile into
 * oracle/_ref/libprt_ref.so.  No reference source is copied into this repository.
 *
 * Tier 1 (as shipped):   ref_render_shipped  -> CPURenderer::Init/ProgressiveRender unmodified
 *                        (depth 20, std::rand, one std::async thread per 16x16 tile).
 * Tier 2 (deterministic): ref_render / ref_trace_pixel -> TraceRay restated verbatim in its
 *                        recursive form (cpu/renderer.cpp:59-103) with a depth parameter, calling the
 *                        reference's compiled Scene::Intersect, MaterialHandle::Emit/Scatter and
 *                        Camera::GetCameraRay.  std::rand() (src/core/math.h:15) is interposed by the
 *                        rand() defined below (the .so is linked -Bsymbolic-functions), which returns
 *                        the 31-bit stream of (seed, pixel, sample, bounce) documented in DESIGN.md.
 */
#define private public          /* Film::m_Accum / m_Weights have no getter (film.h:50,54) */
#include <core/film.h>
#undef private
#include <core/scene.h>
#include <core/camera.h>
#include <core/mesh.h>
#include <backend/cpu/renderer.h>
#include <dlfcn.h>
#include <atomic>
#include <chrono>
#include <cstring>
#include <deque>
#include <memory>
#include <thread>
#include <vector>
#include "../parallelraytracing_b200/host/flatten.h"

/* ---------------------------------------------------------------- interposed rand() */
namespace {
inline uint32_t hash32(uint32_t x) { x ^= x >> 16; x *= 0x21f0aaadu; x ^= x >> 15; x *= 0x735a2d97u; x ^= x >> 15; return x; }
inline uint32_t path_key(uint64_t seed, uint32_t pixel, uint32_t sample) {
    uint32_t k = hash32((uint32_t)(seed >> 32) ^ 0x9e3779b9u);
    k = hash32((uint32_t)seed ^ k); k = hash32(sample ^ k); k = hash32(pixel ^ k); return k;
}
thread_local uint32_t tl_state = 0;
thread_local bool tl_deterministic = false;
thread_local uint64_t tl_draws = 0;
inline void rng_begin_bounce(uint32_t key, uint32_t bounce) { tl_state = hash32(key + 0x9e3779b9u * (bounce + 1u)); tl_deterministic = true; }
unsigned int g_fallback_seed = 12345u;
}  // namespace

extern "C" int rand(void) {
    if (!tl_deterministic) {                     /* tier 1: the real libc rand(), global lock and all */
        static int (*real_rand)(void) = (int (*)(void))dlsym(RTLD_NEXT, "rand");
        return real_rand ? real_rand() : (int)(rand_r(&g_fallback_seed) & 0x7fffffff);
    }
    ++tl_draws;
    tl_state = tl_state * 747796405u + 2891336453u;
    uint32_t w = ((tl_state >> ((tl_state >> 28) + 4u)) ^ tl_state) * 277803737u;
    w = (w >> 22) ^ w;
    return (int)(w >> 1);
}

/* ---------------------------------------------------------------- scene wrapper */
namespace {
struct RefScene {
    std::unique_ptr<Scene> preset;
    MaterialPool mats; ShapePool shapes; std::deque<Triangle> tris; PrimitiveList list;
    const std::vector<Primitive>& prims() const { return preset ? preset->GetPrimitives() : list.GetPrimitives(); }
    void Intersect(const Ray& r, SurfaceInteraction* si) const { if (preset) preset->Intersect(r, si); else list.Intersect(r, si); }
};

const glm::vec3 kSky{0.4f, 0.3f, 0.6f};      /* m_SkyLight, cpu/renderer.h:31 */

/* CPURenderer::TraceRay, cpu/renderer.cpp:59-103, with `bounce` threaded through for the RNG key */
glm::vec3 TraceRay(const RefScene& sc, const Ray& ray, int depth, uint32_t key, uint32_t bounce, uint64_t* rays) {
    if (depth <= 0) return glm::vec3{0.0f};
    glm::vec3 L{0.0f};
    SurfaceInteraction intersect;
    ++*rays;
    sc.Intersect(ray, &intersect);
    if (intersect.HasIntersection) {
        glm::vec3 emittedColor{0.0f};
        intersect.Material.Emit(emittedColor);
        L += emittedColor;
        glm::vec3 attenuation;
        Ray scatteredRay;
        rng_begin_bounce(key, bounce);
        bool scattered = intersect.Material.Scatter(ray, intersect, attenuation, scatteredRay);
        if (scattered) {
            scatteredRay.Normalize();
            L += attenuation * TraceRay(sc, scatteredRay, depth - 1, key, bounce + 1, rays);
        }
    } else {
        L += kSky;
    }
    return L;
}

Camera make_camera(const float* pos, const float* front, int W, int H) {
    return Camera(glm::vec3(pos[0], pos[1], pos[2]), glm::vec3(front[0], front[1], front[2]), (float)W, (float)H, 100.0f);
}
}  // namespace

extern "C" {

void* ref_scene_preset(int preset) {
    RefScene* s = new RefScene();
    s->preset.reset(new Scene((ScenePreset)preset));
    return s;
}

void ref_scene_destroy(void* s) { delete (RefScene*)s; }
int ref_num_prims(void* s) { return (int)((RefScene*)s)->prims().size(); }

/* Custom scene builder: primitives are appended in call order (list order = prim_id). */
void* ref_scene_new() { return new RefScene(); }
static MaterialHandle add_mat(RefScene* s, const prt_material* m) {
    switch (m->type) {
        case PRT_MAT_LAMBERTIAN: return s->mats.AddLambertian(glm::vec3(m->r, m->g, m->b));
        case PRT_MAT_METAL: return s->mats.AddMetal(glm::vec3(m->r, m->g, m->b), m->param);
        case PRT_MAT_DIELECTRIC: return s->mats.AddDielectric(m->param);
        case PRT_MAT_EMISSIVE: return s->mats.AddEmissive(glm::vec3(m->r, m->g, m->b));
        default: return MaterialHandle{};
    }
}
void ref_scene_add_sphere(void* sv, float radius, const float* scale, const float* euler_deg, const float* trans, const prt_material* m) {
    RefScene* s = (RefScene*)sv;
    Primitive p; p.Shape = s->shapes.AddCircle(radius); p.Material = add_mat(s, m);
    p.Transform.Set(glm::vec3(scale[0], scale[1], scale[2]), glm::radians(glm::vec3(euler_deg[0], euler_deg[1], euler_deg[2])), glm::vec3(trans[0], trans[1], trans[2]));
    s->list.AddCircle(p);
}
void ref_scene_add_quad(void* sv, float w, float h, const float* scale, const float* euler_deg, const float* trans, const prt_material* m) {
    RefScene* s = (RefScene*)sv;
    Primitive p; p.Shape = s->shapes.AddQuad(w, h); p.Material = add_mat(s, m);
    p.Transform.Set(glm::vec3(scale[0], scale[1], scale[2]), glm::radians(glm::vec3(euler_deg[0], euler_deg[1], euler_deg[2])), glm::vec3(trans[0], trans[1], trans[2]));
    s->list.AddQuad(p);
}
/* nt triangles (9 floats verts + 9 floats normals each), identity transform, one material */
void ref_scene_add_triangles(void* sv, const float* verts, const float* normals, int nt, const prt_material* m) {
    RefScene* s = (RefScene*)sv;
    MaterialHandle mh = add_mat(s, m);
    for (int64_t t = 0; t < nt; ++t) {
        const float* v = verts + 9 * t; const float* n = normals + 9 * t;
        s->tris.emplace_back(glm::vec3(v[0], v[1], v[2]), glm::vec3(v[3], v[4], v[5]), glm::vec3(v[6], v[7], v[8]),
                             glm::vec3(n[0], n[1], n[2]), glm::vec3(n[3], n[4], n[5]), glm::vec3(n[6], n[7], n[8]),
                             glm::vec2(0.f), glm::vec2(0.f), glm::vec2(0.f));
        Primitive p; p.Shape = MakeShapeHandle(ShapeType::TRIANGLE, &s->tris.back()); p.Material = mh;
        s->list.AddPrimitive(p);
    }
}
/* PLY through the reference's Mesh loader (mesh.cpp:23-184): returns counts, then data */
void* ref_mesh_load(const char* path) { return new Mesh(std::string(path)); }
void ref_mesh_destroy(void* m) { delete (Mesh*)m; }
int ref_mesh_counts(void* m, int* nv, int* nn, int* ni) {
    Mesh* me = (Mesh*)m; *nv = (int)me->GetVertices().size(); *nn = (int)me->GetNormals().size(); *ni = (int)me->GetIndices().size(); return 0;
}
void ref_mesh_data(void* m, float* verts, float* normals, uint32_t* indices) {
    Mesh* me = (Mesh*)m;
    if (verts) std::memcpy(verts, me->GetVertices().data(), me->GetVertices().size() * 12);
    if (normals) std::memcpy(normals, me->GetNormals().data(), me->GetNormals().size() * 12);
    if (indices) std::memcpy(indices, me->GetIndices().data(), me->GetIndices().size() * 4);
}

/* ---- flatten (product host code, parallelraytracing_b200/host/flatten.h) on the reference scene */
int ref_flatten_counts(void* s, int* ns, int* nq, int* nt, int* nm, int* tri_prim_base) {
    prt_host::FlatScene fs = prt_host::FlattenPrimitives(((RefScene*)s)->prims());
    *ns = (int)fs.spheres.size(); *nq = (int)fs.quads.size(); *nt = (int)fs.tri_material.size(); *nm = (int)fs.materials.size();
    *tri_prim_base = fs.tri_prim_base;
    return fs.ok ? 0 : -1;
}
int ref_flatten(void* s, prt_sphere* spheres, prt_quad* quads, float* tv, float* tn, int32_t* tm, prt_material* mats) {
    prt_host::FlatScene fs = prt_host::FlattenPrimitives(((RefScene*)s)->prims());
    if (spheres) std::memcpy(spheres, fs.spheres.data(), fs.spheres.size() * sizeof(prt_sphere));
    if (quads) std::memcpy(quads, fs.quads.data(), fs.quads.size() * sizeof(prt_quad));
    if (tv) std::memcpy(tv, fs.tri_verts.data(), fs.tri_verts.size() * 4);
    if (tn) std::memcpy(tn, fs.tri_normals.data(), fs.tri_normals.size() * 4);
    if (tm) std::memcpy(tm, fs.tri_material.data(), fs.tri_material.size() * 4);
    if (mats) std::memcpy(mats, fs.materials.data(), fs.materials.size() * sizeof(prt_material));
    return fs.ok ? 0 : -1;
}
void ref_camera(const float* pos, const float* front, int W, int H, prt_camera* out) {
    Camera cam = make_camera(pos, front, W, H);
    *out = prt_host::FlattenCamera(cam);
}
/* Camera::GetCameraRay at every pixel centre: W*H*3 floats */
void ref_camera_dirs(const float* pos, const float* front, int W, int H, float* out) {
    Camera cam = make_camera(pos, front, W, H);
    for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
        Ray r = cam.GetCameraRay((float)x + 0.5f, (float)y + 0.5f);
        float* o = out + 3 * ((size_t)y * W + x); o[0] = r.Direction.x; o[1] = r.Direction.y; o[2] = r.Direction.z;
    }
}

/* ---- closest hit through PrimitiveList::Intersect (primitive.cpp:21-59).
 * prim_id is recovered by repeating the per-primitive step with the reference's own functions and
 * matching the winning Position bit-for-bit; second_d (optional) = distance of the closest OTHER
 * primitive hit (inf if none) so tests can mask near-ties. t = sqrt(distance2). */
void ref_trace_batch(void* sv, const prt_ray* rays, int n, prt_hit* out, float* second_d, int n_threads) {
    const RefScene& sc = *(RefScene*)sv;
    const std::vector<Primitive>& prims = sc.prims();
    if (n_threads < 1) n_threads = 1;
    auto work = [&](int tid) {
        for (int i = tid; i < n; i += n_threads) {
            Ray ray; ray.Origin = glm::vec3(rays[i].ox, rays[i].oy, rays[i].oz); ray.Direction = glm::vec3(rays[i].dx, rays[i].dy, rays[i].dz);
            SurfaceInteraction si;
            sc.Intersect(ray, &si);
            prt_hit r; std::memset(&r, 0, sizeof r);
            float best2 = INFINITY, second2 = INFINITY; int best_i = -1;
            if (si.HasIntersection || second_d) {
                for (int k = 0; k < (int)prims.size(); ++k) {
                    const Primitive& p = prims[k];
                    Ray rl; rl.Origin = TransformPoint(p.Transform.GetInvMat(), ray.Origin);
                    rl.Direction = TransformNormal(p.Transform.GetMat(), ray.Direction);
                    SurfaceInteraction s2;
                    if (!p.Shape.Intersect(rl, &s2)) continue;
                    glm::vec3 P = TransformPoint(p.Transform.GetMat(), s2.Position);
                    glm::vec3 dv = ray.Origin - P; float d2 = glm::dot(dv, dv);
                    if (d2 < best2) { second2 = best2; best2 = d2; best_i = k; }
                    else if (d2 < second2) second2 = d2;
                }
            }
            if (!si.HasIntersection) { r.t = INFINITY; r.prim_id = -1; r.material = -1; }
            else {
                glm::vec3 dv = ray.Origin - si.Position;
                r.t = std::sqrt(glm::dot(dv, dv)); r.prim_id = best_i;
                r.px = si.Position.x; r.py = si.Position.y; r.pz = si.Position.z;
                r.nx = si.Normal.x; r.ny = si.Normal.y; r.nz = si.Normal.z;
                r.front = si.IsFrontFace ? 1 : 0; r.material = (int)si.Material.Tag();   /* MatType tag, not an index */
            }
            out[i] = r;
            if (second_d) second_d[i] = std::sqrt(second2);
        }
    };
    if (n_threads == 1) work(0);
    else { std::vector<std::thread> th; for (int t = 0; t < n_threads; ++t) th.emplace_back(work, t); for (auto& t : th) t.join(); }
}

/* ---- tier 2: deterministic render; accum += samples [sample_begin, +spp). Returns #Intersect calls */
uint64_t ref_render(void* sv, const float* pos, const float* front, int W, int H, int spp, int depth, uint64_t seed,
                    int sample_begin, float* accum, int n_threads, int y_begin, int y_end) {
    const RefScene& sc = *(RefScene*)sv;
    Camera cam = make_camera(pos, front, W, H);
    if (n_threads < 1) n_threads = 1;
    if (y_end <= 0 || y_end > H) y_end = H;
    if (y_begin < 0) y_begin = 0;
    std::vector<uint64_t> rays(n_threads, 0);
    std::atomic<int> next_row(y_begin);
    auto work = [&](int tid) {
        for (;;) {
            int y = next_row.fetch_add(1);
            if (y >= y_end) break;
            for (int x = 0; x < W; ++x) {
                uint32_t pixel = (uint32_t)(y * W + x);
                Ray ray = cam.GetCameraRay((float)x + 0.5f, (float)y + 0.5f);          /* cpu/renderer.cpp:45 */
                float* a = accum + 3 * (size_t)pixel;
                for (int s = 0; s < spp; ++s) {
                    glm::vec3 L = TraceRay(sc, ray, depth, path_key(seed, pixel, (uint32_t)(sample_begin + s)), 0, &rays[tid]);
                    a[0] += L.r; a[1] += L.g; a[2] += L.b;                               /* Film::AddSample, film.cu:37-55 */
                }
            }
        }
        tl_deterministic = false;
    };
    if (n_threads == 1) work(0);
    else { std::vector<std::thread> th; for (int t = 0; t < n_threads; ++t) th.emplace_back(work, t); for (auto& t : th) t.join(); }
    uint64_t total = 0; for (uint64_t r : rays) total += r;
    return total;
}

void ref_trace_pixel(void* sv, const float* pos, const float* front, int W, int H, int x, int y, int sample, int depth, uint64_t seed, float* rgb) {
    const RefScene& sc = *(RefScene*)sv;
    Camera cam = make_camera(pos, front, W, H);
    uint64_t rays = 0;
    Ray ray = cam.GetCameraRay((float)x + 0.5f, (float)y + 0.5f);
    glm::vec3 L = TraceRay(sc, ray, depth, path_key(seed, (uint32_t)(y * W + x), (uint32_t)sample), 0, &rays);
    tl_deterministic = false;
    rgb[0] = L.r; rgb[1] = L.g; rgb[2] = L.b;
}

/* ---- tier 1: the reference backend exactly as shipped (preset scenes only). Returns seconds for
 * `spp` ProgressiveRender() calls; accum_out (W*H*3, optional) = Film's float accumulator. */
double ref_render_shipped(int preset, const float* pos, const float* front, int W, int H, int spp, float* accum_out) {
    Film film((uint32_t)W, (uint32_t)H);
    Scene scene((ScenePreset)preset);
    Camera cam = make_camera(pos, front, W, H);
    CPURenderer r;
    r.Init(film, scene, cam);
    auto t0 = std::chrono::steady_clock::now();
    for (int s = 0; s < spp; ++s) r.ProgressiveRender();
    auto t1 = std::chrono::steady_clock::now();
    if (accum_out) std::memcpy(accum_out, film.m_Accum.data(), (size_t)W * H * 3 * sizeof(float));
    return std::chrono::duration<double>(t1 - t0).count();
}

/* ---- Film::AddSampleBuffer + UpdateDisplay (film.cu:57-77, 134-194) on a sum-of-samples buffer */
void ref_film_display(const float* accum_sum, int W, int H, int spp, float exposure, float gamma, uint8_t* rgba) {
    Film film((uint32_t)W, (uint32_t)H);
    std::vector<float> mean((size_t)W * H * 3);
    for (size_t i = 0; i < mean.size(); ++i) mean[i] = accum_sum[i];
    /* Accum += rgb*weight, Weights += weight: pass the sum with weight 1, then fix the weights */
    film.AddSampleBuffer(mean.data(), 1.0f);
    for (auto& w : film.m_Weights) w = (float)spp;
    film.UpdateDisplay(exposure, gamma);
    std::memcpy(rgba, film.GetDisplayData(), (size_t)W * H * 4);
}

/* ---- unit hooks: one Scatter call with the (seed,pixel,sample,bounce) stream */
int ref_scatter(const prt_material* m, const float* in_dir, const float* pos, const float* normal, int front,
                uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, float* atten, float* out_dir, float* emit) {
    RefScene tmp; MaterialHandle mh = add_mat(&tmp, m);
    SurfaceInteraction si; si.Position = glm::vec3(pos[0], pos[1], pos[2]); si.Normal = glm::vec3(normal[0], normal[1], normal[2]);
    si.HasIntersection = true; si.IsFrontFace = front != 0; si.Material = mh;
    Ray in; in.Origin = glm::vec3(0.f); in.Direction = glm::vec3(in_dir[0], in_dir[1], in_dir[2]);
    glm::vec3 e(0.f), a(0.f); Ray out;
    mh.Emit(e);
    rng_begin_bounce(path_key(seed, pixel, sample), bounce);
    bool ok = mh.Scatter(in, si, a, out);
    tl_deterministic = false;
    if (ok) out.Normalize();
    emit[0] = e.x; emit[1] = e.y; emit[2] = e.z; atten[0] = a.x; atten[1] = a.y; atten[2] = a.z;
    out_dir[0] = out.Direction.x; out_dir[1] = out.Direction.y; out_dir[2] = out.Direction.z;
    return ok ? 1 : 0;
}

}  /* extern "C" */
