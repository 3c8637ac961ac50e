/*
 * shim_driver.cpp — TEST INFRASTRUCTURE (oracle/): the reference's own viewer call pattern
 * (main.cpp:138-170 Film/Scene/Camera/Init, :506-532 SetCamera + ProgressiveRender + UpdateDisplayGPU),
 * headless, driving the drop-in backend `B200Renderer : public Renderer`
 * (parallelraytracing_b200/host/b200_renderer.{h,cpp}) through the reference's unmodified
 * Renderer / Film / Scene / Camera classes. Linked with the reference objects built by oracle/Makefile
 * and with libprt_b200.so. Dumps Film's device accumulators so tests can compare them with the C-ABI path.
 *   shim_driver <preset> <W> <H> <calls> <spp_per_call> <depth> <seed> <out.bin> [<mesh.ply> <mat_type> <r> <g> <b> <param> <tx> <ty> <tz>]...
 * Every trailing 9-tuple loads a PLY with the reference's own Mesh class (mesh.cpp:23-184) and registers it with
 * B200Renderer::AddMesh (the reference Scene cannot hold meshes).
 */
#define private public          /* Film::d_Accum / d_Weights have no getter (film.h:50-55) */
#include <core/film.h>
#undef private
#include <core/scene.h>
#include <core/camera.h>
#include <core/renderer.h>
#include <core/mesh.h>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>
#include "../parallelraytracing_b200/host/b200_renderer.h"

int main(int argc, char** argv)
{
    if (argc < 9) { std::fprintf(stderr, "usage: %s preset W H calls spp_per_call depth seed out.bin\n", argv[0]); return 2; }
    const int preset = std::atoi(argv[1]), W = std::atoi(argv[2]), H = std::atoi(argv[3]), calls = std::atoi(argv[4]);
    const int spp = std::atoi(argv[5]), depth = std::atoi(argv[6]);
    const unsigned long long seed = std::strtoull(argv[7], nullptr, 10);

    Film film{(uint32_t)W, (uint32_t)H};                                  // main.cpp:138
    Scene scene{(ScenePreset)preset};                                     // main.cpp:140
    glm::vec3 pos{5.0f, 5.0f, 8.0f};
    Camera camera{pos, glm::normalize(glm::vec3{0.0f} - pos), (float)W, (float)H, 100.0f};   // main.cpp:142-150

    auto backend = std::make_shared<B200Renderer>();
    backend->SetMaxDepth(depth); backend->SetSamplesPerCall(spp); backend->SetSeed(seed);
    if (const char* devs = std::getenv("PRT_SHIM_DEVICES")) {              // e.g. "0,1": one Renderer over several GPUs (SetDevices)
        std::vector<int> ids;
        for (const char* p = devs; *p;) { ids.push_back(std::atoi(p)); while (*p && *p != ',') ++p; if (*p == ',') ++p; }
        backend->SetDevices(ids);
    }
    std::vector<std::unique_ptr<Mesh>> meshes;
    for (int a = 9; a + 8 < argc; a += 9) {
        meshes.emplace_back(new Mesh(argv[a]));
        prt_material m{};
        m.type = std::atoi(argv[a + 1]); m.r = (float)std::atof(argv[a + 2]); m.g = (float)std::atof(argv[a + 3]);
        m.b = (float)std::atof(argv[a + 4]); m.param = (float)std::atof(argv[a + 5]);
        backend->AddMesh(*meshes.back(), m, (float)std::atof(argv[a + 6]), (float)std::atof(argv[a + 7]), (float)std::atof(argv[a + 8]));
    }
    std::shared_ptr<Renderer> renderer = backend;                          // held as the abstract type, main.cpp:152-163
    renderer->Init(film, scene, camera);                                   // main.cpp:165-166
    film.Clear();                                                          // main.cpp:168
    if (!backend->Ok()) { std::fprintf(stderr, "B200Renderer::Init failed\n"); return 1; }
    renderer->SetCamera(camera); film.Clear();                             // main.cpp:506-510
    for (int i = 0; i < calls; ++i) {
        renderer->ProgressiveRender();                                     // main.cpp:513
        film.UpdateDisplayGPU();                                           // main.cpp:516
    }
    cudaDeviceSynchronize();
    const size_t np = (size_t)W * H;
    std::vector<float> accum(np * 3), weights(np);
    std::vector<unsigned char> disp(np * 4);
    cudaMemcpy(accum.data(), film.d_Accum, np * 12, cudaMemcpyDeviceToHost);
    cudaMemcpy(weights.data(), film.d_Weights, np * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(disp.data(), film.GetDisplayDataGPU(), np * 4, cudaMemcpyDeviceToHost);
    FILE* f = std::fopen(argv[8], "wb");
    if (!f) return 3;
    std::fwrite(accum.data(), 4, accum.size(), f);
    std::fwrite(weights.data(), 4, weights.size(), f);
    std::fwrite(disp.data(), 1, disp.size(), f);
    std::fclose(f);
    std::printf("shim ok: %d calls x %d spp, %dx%d\n", calls, spp, W, H);
    return 0;
}
