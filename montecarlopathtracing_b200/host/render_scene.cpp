// render_scene(path, filename, spp): the reference's entry point (MTPC/MTPC.cpp:35-68) re-plumbed
// onto the C ABI: load -> upload -> GPU LBVH build -> GPU render -> quantise -> PNG.
#include <cstdio>
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
}  // namespace

bool render_scene(std::string path, std::string filename, int N_ray_per_pixel) {
    int rc = render_to_png(path, filename, N_ray_per_pixel, "../result/", 0);
    if (rc) std::fprintf(stderr, "render_scene failed (%d): %s\n", rc, mcpt_last_error());
    return rc == 0;
}
bool render_scene(std::string filename, int N_ray_per_pixel) { return render_scene("../scene/", filename, N_ray_per_pixel); }

extern "C" MCPT_API int mcpt_render_scene(const char* path, const char* name, int spp, const char* out_dir, int device) {
    if (!path || !name || spp <= 0) { mcpt_set_error("mcpt_render_scene: bad argument"); return MCPT_ERR_ARG; }
    return render_to_png(path, name, spp, out_dir ? out_dir : "../result/", device);
}
