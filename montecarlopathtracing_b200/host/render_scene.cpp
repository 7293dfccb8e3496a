// render_scene(path, filename, spp): the reference's entry point (MTPC/MTPC.cpp:35-68) re-plumbed
// onto the C ABI: load -> upload -> GPU LBVH build -> GPU render -> quantise -> PNG.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include "scene_loader.h"

void mcpt_set_error(const char* msg);

namespace {
int render_to_png(const std::string& path, const std::string& name, int spp, const std::string& out_dir, int device) {
    mcpt_ctx* ctx = nullptr;
    mcpt_camera camera{};
    const mcpt_camera* cam = &camera;
    mcpt_stats st{};
    float build_ms = 0, ingest_ms = 0;
    int rc = mcpt_create(device, &ctx);
    // .mtl / .camera on the host, the .obj parsed and flattened on the GPU (bit-identical to mcpt_scene_load + mcpt_upload_scene)
    if (!rc) rc = mcpt_scene_load_gpu(ctx, path.c_str(), name.c_str(), &camera, &ingest_ms);
    if (!rc) rc = mcpt_build_bvh(ctx, &build_ms);
    const std::string file = out_dir + name + "-SPP" + std::to_string(spp) + ".png";  // MTPC/MTPC.cpp:19-20,64-65
    // integrator + imshow/svpng on the device: only the finished PNG crosses PCIe
    if (!rc) rc = mcpt_render_png(ctx, cam, 0, 0, spp, MCPT_MODE_WAVEFRONT, 0, file.c_str(), &st);
    if (!rc) {
        std::printf("Phase 0 (GPU OBJ ingest) = %.3f ms\nPhase 1 (GPU LBVH build) = %.3f ms\nPhase 2 (path tracing) = %.3f ms, %.1f Msamples/s, %.1f Mrays/s\n",
                    ingest_ms, build_ms, st.render_ms, st.samples / st.render_ms * 1e-3, st.rays / st.render_ms * 1e-3);
    }
    if (ctx) mcpt_destroy(ctx);
    return rc;
}

// The same over n_gpus GPUs of this process (SPP split, one NCCL exchange, fixed-order sum: the same PNG bytes as one GPU).
int render_to_png_multi(const std::string& path, const std::string& name, int spp, const std::string& out_dir, int n_gpus) {
    mcpt_multi* m = nullptr;
    mcpt_camera camera{};
    mcpt_stats st{};
    float build_ms = 0;
    int rc = mcpt_multi_create(nullptr, n_gpus, &m);
    if (!rc) rc = mcpt_multi_load_scene_gpu(m, path.c_str(), name.c_str(), &camera);
    if (!rc) rc = mcpt_multi_build_bvh(m, &build_ms);
    const std::string file = out_dir + name + "-SPP" + std::to_string(spp) + ".png";
    if (!rc) rc = mcpt_multi_render_png(m, &camera, 0, 0, spp, MCPT_MODE_WAVEFRONT, 0, file.c_str(), &st);
    if (!rc)
        std::printf("Phase 1 (GPU LBVH build, %d GPUs) = %.3f ms\nPhase 2 (path tracing) = %.3f ms + %.3f ms exchange, %.1f Msamples/s, %.1f Mrays/s\n",
                    n_gpus, build_ms, st.render_ms, st.exchange_ms, st.samples / (st.render_ms + st.exchange_ms) * 1e-3,
                    st.rays / (st.render_ms + st.exchange_ms) * 1e-3);
    if (m) mcpt_multi_destroy(m);
    return rc;
}
// How many GPUs render_scene uses when nobody says: every visible one (a power of two <= 8) once the job is large enough to
// pay for bringing up the communicator (~1 s for 8 GPUs): at least 2^28 samples per GPU, ~0.15 s of rendering each.
// MCPT_GPUS overrides.
int default_gpu_count(const std::string& path, const std::string& name, int spp) {
    if (const char* v = std::getenv("MCPT_GPUS")) { const int n = std::atoi(v); if (n >= 1) return n; }
    int visible = 0;
    {   // device count through the C ABI's own probe: a context on device k exists iff k < count
        mcpt_ctx* c = nullptr;
        while (visible < 8 && mcpt_create(visible, &c) == MCPT_OK) { mcpt_destroy(c); c = nullptr; visible++; }
    }
    if (visible <= 1) return 1;
    mcpt_host_scene* hs = nullptr;   // only the camera's resolution is needed
    long long pixels = 1024LL * 1024LL;
    if (mcpt_scene_load(path.c_str(), name.c_str(), &hs) == MCPT_OK) {
        const mcpt_camera* cam = mcpt_scene_get_camera(hs);
        pixels = (long long)cam->width * cam->height;
        mcpt_scene_free(hs);
    }
    int n = 1;
    while (n * 2 <= visible && n * 2 <= 8 && pixels * spp / (n * 2) >= (1LL << 28)) n *= 2;
    return n;
}
}  // namespace

bool render_scene(std::string path, std::string filename, int N_ray_per_pixel) {
    const int n_gpus = default_gpu_count(path, filename, N_ray_per_pixel);
    int rc = n_gpus > 1 ? render_to_png_multi(path, filename, N_ray_per_pixel, "../result/", n_gpus)
                        : render_to_png(path, filename, N_ray_per_pixel, "../result/", 0);
    if (rc) std::fprintf(stderr, "render_scene failed (%d): %s\n", rc, mcpt_last_error());
    return rc == 0;
}
bool render_scene(std::string filename, int N_ray_per_pixel) { return render_scene("../scene/", filename, N_ray_per_pixel); }

extern "C" MCPT_API int mcpt_render_scene(const char* path, const char* name, int spp, const char* out_dir, int device) {
    if (!path || !name || spp <= 0) { mcpt_set_error("mcpt_render_scene: bad argument"); return MCPT_ERR_ARG; }
    return render_to_png(path, name, spp, out_dir ? out_dir : "../result/", device);
}

extern "C" MCPT_API int mcpt_render_scene_multi(const char* path, const char* name, int spp, const char* out_dir, int n_gpus) {
    if (!path || !name || spp <= 0 || n_gpus < 0) { mcpt_set_error("mcpt_render_scene_multi: bad argument"); return MCPT_ERR_ARG; }
    if (n_gpus == 0) {
        int visible = 0;
        mcpt_ctx* c = nullptr;
        while (visible < 8 && mcpt_create(visible, &c) == MCPT_OK) { mcpt_destroy(c); c = nullptr; visible++; }
        n_gpus = 1;
        while (n_gpus * 2 <= visible) n_gpus *= 2;
    }
    return render_to_png_multi(path, name, spp, out_dir ? out_dir : "../result/", n_gpus);
}
