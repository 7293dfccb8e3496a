// mcpt_cli — the reference's main() (MTPC/MTPC.cpp:71-91) with its hard-coded scene name and spp turned into
// arguments.  Defaults reproduce the reference: render_scene("../scene/", "cornell-box", 25) -> ../result/.
//   mcpt_cli [--scene-dir DIR/] [--name NAME] [--spp N] [--width W --height H] [--mode wavefront|megakernel]
//            [--out DIR/] [--device I] [--seed S] [--gpus N]
// --gpus N (1, 2, 4 or 8): SPP split over N GPUs of this process (mcpt_multi_*: ncclCommInitAll, one exchange, fixed-order
// sum); the PNG is the same file, byte for byte, as with --gpus 1.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/mcpt.h"

static int fail(const char* what, int rc) {
    std::fprintf(stderr, "mcpt_cli: %s failed (%d): %s\n", what, rc, mcpt_last_error());
    return 1;
}

int main(int argc, char** argv) {
    std::string dir = "../scene/", name = "cornell-box", out = "../result/";
    int spp = 25, width = 0, height = 0, mode = MCPT_MODE_WAVEFRONT, device = 0, gpus = 1;
    unsigned long long seed = 0;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&]() -> const char* { if (i + 1 >= argc) { std::fprintf(stderr, "mcpt_cli: %s needs a value\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--scene-dir") dir = next();
        else if (a == "--name") name = next();
        else if (a == "--spp") spp = std::atoi(next());
        else if (a == "--width") width = std::atoi(next());
        else if (a == "--height") height = std::atoi(next());
        else if (a == "--out") out = next();
        else if (a == "--device") device = std::atoi(next());
        else if (a == "--gpus") gpus = std::atoi(next());
        else if (a == "--seed") seed = std::strtoull(next(), nullptr, 10);
        else if (a == "--mode") { const std::string m = next(); mode = (m == "megakernel") ? MCPT_MODE_MEGAKERNEL : MCPT_MODE_WAVEFRONT; }
        else { std::fprintf(stderr, "usage: mcpt_cli [--scene-dir DIR/] [--name NAME] [--spp N] [--width W --height H] "
                                    "[--mode wavefront|megakernel] [--out DIR/] [--device I] [--seed S] [--gpus N]\n"); return 2; }
    }
    mcpt_host_scene* hs = nullptr;
    int rc = mcpt_scene_load(dir.c_str(), name.c_str(), &hs);
    if (rc) return fail("mcpt_scene_load", rc);
    const mcpt_camera* cam = mcpt_scene_get_camera(hs);
    const int w = width > 0 ? width : cam->width, h = height > 0 ? height : cam->height;
    const std::string file = out + name + "-SPP" + std::to_string(spp) + ".png";   // MTPC/MTPC.cpp:19-20,64-65
    if (gpus > 1) {
        mcpt_multi* m = nullptr;
        if ((rc = mcpt_multi_create(nullptr, gpus, &m))) return fail("mcpt_multi_create", rc);
        mcpt_multi_set_seed(m, seed);
        if ((rc = mcpt_multi_upload_scene(m, mcpt_scene_get_desc(hs)))) return fail("mcpt_multi_upload_scene", rc);
        float bms = 0.f;
        if ((rc = mcpt_multi_build_bvh(m, &bms))) return fail("mcpt_multi_build_bvh", rc);
        std::vector<float> img((size_t)w * h * 3);
        mcpt_stats ms{};
        if ((rc = mcpt_multi_render(m, cam, w, h, spp, mode, 0, img.data(), &ms))) return fail("mcpt_multi_render", rc);
        std::printf("Phase 1 (GPU LBVH build, %d GPUs) = %.3f ms\nPhase 2 (path tracing, %s) = %.3f ms + %.3f ms exchange: %.1f Msamples/s, %.1f Mrays/s\n",
                    gpus, bms, mode == MCPT_MODE_WAVEFRONT ? "wavefront" : "megakernel", ms.render_ms, ms.exchange_ms,
                    ms.samples / (ms.render_ms + ms.exchange_ms) * 1e-3, ms.rays / (ms.render_ms + ms.exchange_ms) * 1e-3);
        if ((rc = mcpt_write_png(file.c_str(), img.data(), w, h))) return fail("mcpt_write_png", rc);
        std::printf("wrote %s\n", file.c_str());
        mcpt_multi_destroy(m);
        mcpt_scene_free(hs);
        return 0;
    }
    mcpt_ctx* ctx = nullptr;
    if ((rc = mcpt_create(device, &ctx))) return fail("mcpt_create", rc);
    mcpt_set_seed(ctx, seed);
    if ((rc = mcpt_upload_scene(ctx, mcpt_scene_get_desc(hs)))) return fail("mcpt_upload_scene", rc);
    float build_ms = 0.f;
    if ((rc = mcpt_build_bvh(ctx, &build_ms))) return fail("mcpt_build_bvh", rc);
    std::vector<float> rgb((size_t)w * h * 3);
    mcpt_stats st{};
    if ((rc = mcpt_render(ctx, cam, w, h, spp, mode, 0, rgb.data(), &st))) return fail("mcpt_render", rc);
    std::printf("Phase 1 (GPU LBVH build) = %.3f ms\nPhase 2 (path tracing, %s) = %.3f ms: %.1f Msamples/s, %.1f Mrays/s\n", build_ms,
                mode == MCPT_MODE_WAVEFRONT ? "wavefront" : "megakernel", st.render_ms, st.samples / st.render_ms * 1e-3,
                st.rays / st.render_ms * 1e-3);
    if ((rc = mcpt_write_png(file.c_str(), rgb.data(), w, h))) return fail("mcpt_write_png", rc);
    std::printf("wrote %s\n", file.c_str());
    mcpt_destroy(ctx);
    mcpt_scene_free(hs);
    return 0;
}
