// b200_renderer.h — `B200Renderer : public Renderer`: the drop-in backend.
// Implements the reference's 3-method backend interface (src/core/renderer.h:8-16) on top of
// the C ABI in include/prt.h, with the call pattern main.cpp uses (Init once :165-166,
// SetCamera on camera change :506-510, ProgressiveRender once per frame :513, then the caller
// runs film.UpdateDisplayGPU()). Compiled against the unmodified reference headers.
#pragma once
#include <core/renderer.h>
#include <core/mesh.h>
#include <vector>
#include "prt.h"

class B200Renderer : public Renderer
{
public:
    B200Renderer() = default;
    ~B200Renderer() override;

    void Init(Film& film, const Scene& scene, const Camera& camera) override;
    void ProgressiveRender() override;
    void SetCamera(const Camera& camera) override;

    // Extras (the reference hard-codes these: depth 20 at cpu/renderer.h:34, 1 spp per call).
    void SetMaxDepth(int depth) { m_Depth = depth; }
    void SetSamplesPerCall(int spp) { m_SamplesPerCall = spp < 1 ? 1 : spp; }
    void SetSeed(uint64_t seed) { m_Seed = seed; }
    void SetDevice(int device) { m_Device = device; m_Devices.clear(); }
    // Several GPUs of this process behind the one Renderer: prt_create_multi - the scene and BVH are replicated,
    // every ProgressiveRender splits its samples over the devices and one NCCL sum-reduce lands the sample buffer on
    // devices[0], which must be the device the caller's Film lives on (Film::AddSampleBufferGPU reads it there). Before Init.
    void SetDevices(const std::vector<int>& devices) { m_Devices = devices; if (!devices.empty()) m_Device = devices[0]; }
    // Mesh -> scene glue: the reference Scene cannot hold a Mesh; meshes registered here BEFORE Init are appended
    // to the flattened scene as world-space triangles (after the Scene's own primitives). The Mesh must outlive Init.
    void AddMesh(const Mesh& mesh, const prt_material& material, float tx = 0.f, float ty = 0.f, float tz = 0.f)
    { m_Meshes.push_back({&mesh, material, {tx, ty, tz}}); }
    prt_ctx* Context() const { return m_Ctx; }
    bool Ok() const { return m_Ok; }

private:
    Film* m_Film = nullptr;
    const Scene* m_Scene = nullptr;
    const Camera* m_Camera = nullptr;
    prt_ctx* m_Ctx = nullptr;
    bool m_Ok = false;
    void* m_FilmDone = nullptr;    // cudaEvent_t recorded behind Film::AddSampleBufferGPU on the legacy stream
    int m_Device = 0;
    std::vector<int> m_Devices;    // empty: one device (m_Device)
    int m_Depth = 20;              // m_Depth of the reference backends
    int m_SamplesPerCall = 1;
    int m_NextSample = 0;
    uint64_t m_Seed = 1;
    struct MeshRef { const Mesh* mesh; prt_material material; float offset[3]; };
    std::vector<MeshRef> m_Meshes;
};
