/*
 * ref_gpu_driver.cpp — TEST / MEASUREMENT INFRASTRUCTURE (oracle/), SURVEY 8f-3: the reference's own CUDA backends
 * (src/backend/cuda_wavefront/renderer.cu, src/backend/cuda_megakernel/renderer.cu — compiled UNMODIFIED for sm_100
 * where they lie, objects in oracle/_ref/ only) and the drop-in `B200Renderer`, all held as `std::shared_ptr<Renderer>`
 * and driven by the viewer's call pattern (main.cpp:138-170, 506-532) without the GL window, on the same GPU, so
 * that the two implementations are timed side by side through the SAME interface:
 *     Film film(W,H); Scene scene(preset); Camera camera((5,5,8) -> origin);
 *     renderer->Init(film, scene, camera); film.Clear(); renderer->SetCamera(camera); film.Clear();
 *     warm-up calls; then `frames` x { renderer->ProgressiveRender(); }   (+ film.UpdateDisplayGPU() with `display`=1)
 * The reference backends use their own literals: depth 20, 1 spp per call (cuda_wavefront/renderer.cu:407,
 * cuda_megakernel/renderer.cu:86). B200Renderer is set to depth 20 and `spp_per_call` (1 = the reference's granularity).
 *   ref_gpu_driver <wavefront|megakernel|b200> <preset> <W> <H> <warmup> <frames> <spp_per_call> <display 0|1> [out.bin]
 * Prints one JSON line: wall ms per call (chrono around the loop, cudaDeviceSynchronize on both sides), spp/s, mean
 * radiance of the Film (sanity: the backends must agree on the image mean). out.bin = Film accum + weights (floats).
 */
#define private public          /* Film::d_Accum / d_Weights have no getter (film.h:50-55) */
#include <core/film.h>
#undef private
#include <core/scene.h>
#include <core/camera.h>
#include <core/renderer.h>
#include <backend/cuda_wavefront/renderer.h>
#include <backend/cuda_megakernel/renderer.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>
#include "../parallelraytracing_b200/host/b200_renderer.h"

int main(int argc, char** argv)
{
    if (argc < 9) {
        std::fprintf(stderr, "usage: %s wavefront|megakernel|b200 preset W H warmup frames spp_per_call display [out.bin]\n", argv[0]);
        return 2;
    }
    const char* which = argv[1];
    const int preset = std::atoi(argv[2]), W = std::atoi(argv[3]), H = std::atoi(argv[4]);
    const int warmup = std::atoi(argv[5]), frames = std::atoi(argv[6]), spp = std::atoi(argv[7]), display = std::atoi(argv[8]);

    Film film{(uint32_t)W, (uint32_t)H};
    Scene scene{(ScenePreset)preset};
    glm::vec3 pos{5.0f, 5.0f, 8.0f};
    Camera camera{pos, glm::normalize(glm::vec3{0.0f} - pos), (float)W, (float)H, 100.0f};

    std::shared_ptr<Renderer> renderer;
    int spp_per_call = 1;
    if (!std::strcmp(which, "wavefront")) renderer = std::make_shared<CudaWavefrontRenderer>();
    else if (!std::strcmp(which, "megakernel")) renderer = std::make_shared<CudaMegakernelRenderer>();
    else if (!std::strcmp(which, "b200")) {
        auto b = std::make_shared<B200Renderer>();
        b->SetMaxDepth(20); b->SetSamplesPerCall(spp); b->SetSeed(1);
        spp_per_call = spp < 1 ? 1 : spp;
        renderer = b;
    } else { std::fprintf(stderr, "unknown backend %s\n", which); return 2; }

    renderer->Init(film, scene, camera);
    film.Clear();
    renderer->SetCamera(camera); film.Clear();
    for (int i = 0; i < warmup; ++i) { renderer->ProgressiveRender(); if (display) film.UpdateDisplayGPU(); }
    cudaDeviceSynchronize();
    film.Clear();
    cudaDeviceSynchronize();
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < frames; ++i) { renderer->ProgressiveRender(); if (display) film.UpdateDisplayGPU(); }
    cudaDeviceSynchronize();
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    const cudaError_t err = cudaGetLastError();

    const size_t np = (size_t)W * H;
    std::vector<float> accum(np * 3), weights(np);
    cudaMemcpy(accum.data(), film.d_Accum, np * 12, cudaMemcpyDeviceToHost);
    cudaMemcpy(weights.data(), film.d_Weights, np * 4, cudaMemcpyDeviceToHost);
    double sum = 0.0, wsum = 0.0;
    for (size_t i = 0; i < np * 3; ++i) sum += accum[i];
    for (size_t i = 0; i < np; ++i) wsum += weights[i];
    const double mean = wsum > 0 ? sum / (3.0 * wsum) : 0.0;
    if (argc > 9) {
        FILE* f = std::fopen(argv[9], "wb");
        if (!f) return 3;
        std::fwrite(accum.data(), 4, accum.size(), f);
        std::fwrite(weights.data(), 4, weights.size(), f);
        std::fclose(f);
    }
    std::printf("{\"backend\": \"%s\", \"preset\": %d, \"width\": %d, \"height\": %d, \"prims\": %zu, \"depth\": 20, \"frames\": %d, "
                "\"spp_per_call\": %d, \"display\": %d, \"ms_per_call\": %.4f, \"spp_per_s\": %.3f, \"mean_radiance\": %.6f, "
                "\"weight_per_pixel\": %.3f, \"cuda_error\": \"%s\"}\n",
                which, preset, W, H, scene.GetPrimitives().size(), frames, spp_per_call, display, ms / frames,
                1e3 * frames * spp_per_call / ms, mean, wsum / np, cudaGetErrorString(err));
    return err == cudaSuccess ? 0 : 1;
}
