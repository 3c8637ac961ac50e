// b200_renderer.cpp — see b200_renderer.h. Error behaviour mirrors the reference's GPU
// backends: failures are logged and the renderer becomes a no-op
// (optix/renderer.cpp:1006-1014, cuda_wavefront/renderer.cu:374-391).
#include "b200_renderer.h"
#include "flatten.h"
#include <cstdio>
#include <cuda_runtime_api.h>

B200Renderer::~B200Renderer() { prt_destroy(m_Ctx); if (m_FilmDone) cudaEventDestroy((cudaEvent_t)m_FilmDone); }

void B200Renderer::Init(Film& film, const Scene& scene, const Camera& camera)
{
    m_Film = &film; m_Scene = &scene; m_Ok = false;
    if (!m_Ctx && (m_Devices.size() > 1 ? prt_create_multi(&m_Ctx, m_Devices.data(), (int)m_Devices.size())
                                        : prt_create(&m_Ctx, m_Device)) != PRT_OK) {
        std::fprintf(stderr, "[B200Renderer] %s\n", prt_last_error(nullptr));
        return;
    }
    prt_host::FlatScene fs = prt_host::FlattenScene(scene);
    if (!fs.ok) { std::fprintf(stderr, "[B200Renderer] %s\n", fs.error.c_str()); return; }
    if (!m_Meshes.empty()) {
        if (fs.tri_material.empty()) fs.tri_prim_base = (int)scene.GetPrimitives().size();
        else if (fs.tri_prim_base + (int)fs.tri_material.size() != (int)scene.GetPrimitives().size()) {
            std::fprintf(stderr, "[B200Renderer] AddMesh: the scene's own triangles must end its primitive list\n"); return;
        }
        for (const MeshRef& m : m_Meshes)
            prt_host::AppendMesh(fs, *m.mesh, m.material, glm::vec3(m.offset[0], m.offset[1], m.offset[2]));
    }
    const float sky[3] = {0.4f, 0.3f, 0.6f};              // m_SkyLight, cpu/renderer.h:31
    int nt = (int)fs.tri_material.size();
    // scene upload + BVH build in one call: the build runs on the positions while the normals are still crossing PCIe
    if (prt_set_scene_build(m_Ctx, fs.spheres.data(), (int)fs.spheres.size(), fs.quads.data(), (int)fs.quads.size(),
                            fs.tri_verts.data(), fs.tri_normals.data(), fs.tri_material.data(), -1, nt, fs.tri_prim_base,
                            fs.materials.data(), (int)fs.materials.size(), sky) != PRT_OK) {
        std::fprintf(stderr, "[B200Renderer] %s\n", prt_last_error(m_Ctx));
        return;
    }
    m_Ok = true;
    SetCamera(camera);
}

void B200Renderer::SetCamera(const Camera& camera)
{
    m_Camera = &camera;
    if (!m_Ctx || !m_Ok) return;
    prt_camera k = prt_host::FlattenCamera(camera);
    if (prt_set_camera(m_Ctx, &k) != PRT_OK) { std::fprintf(stderr, "[B200Renderer] %s\n", prt_last_error(m_Ctx)); m_Ok = false; }
}

void B200Renderer::ProgressiveRender()
{
    if (!m_Ctx || !m_Ok || !m_Film || !m_Camera) return;
    if (m_Film->GetWidth() != (uint32_t)m_Camera->GetWidth() || m_Film->GetHeight() != (uint32_t)m_Camera->GetHeight()) return;
    const int k = m_SamplesPerCall;
    // Film::AddSampleBufferGPU of the previous frame ran on the legacy default stream and returned without waiting
    // (film.cu:90-99); the context's streams are non-blocking, so the sample buffer must not be rewritten before it is read:
    // the render stream waits ON THE DEVICE for the event recorded behind that kernel (no host synchronisation)
    if (m_FilmDone) prt_wait_event(m_Ctx, m_FilmDone);
    if (prt_render_fresh(m_Ctx, k, m_Depth, m_Seed, m_NextSample) != PRT_OK) {          // clear + render; 1 spp writes the buffer directly
        std::fprintf(stderr, "[B200Renderer] %s\n", prt_last_error(m_Ctx));
        return;
    }
    m_NextSample += k;
    if (k > 1) prt_scale_accum(m_Ctx, 1.0f / (float)k);    // Film adds rgb*weight: hand it the mean with weight k
    m_Film->AddSampleBufferGPU(prt_accum_device_ptr(m_Ctx), (float)k);   // film.cu:90-99
    if (!m_FilmDone) cudaEventCreateWithFlags((cudaEvent_t*)&m_FilmDone, cudaEventDisableTiming);
    if (m_FilmDone) cudaEventRecord((cudaEvent_t)m_FilmDone, cudaStreamLegacy);
}
