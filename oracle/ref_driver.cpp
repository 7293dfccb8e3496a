// TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// Driver/exporter linked against the UNMODIFIED reference objects (compiled by oracle/build_ref.sh
// from /root/reference/MTPC/*.cpp where they lie).  It only calls the reference's own functions:
//   render_scene            MTPC/MTPC.cpp:35
//   scene_data::read_scene  MTPC/sceneManagement.cpp:264
//   compare                 MTPC/sceneManagement.cpp:311
//   BVH::BVH                MTPC/BVH.cpp:37
//   ray_intersect           MTPC/pathTracing.cpp:382
//   generateImg             MTPC/pathTracing.cpp:274
// The one deliberate difference from render_scene: std::stable_sort instead of std::sort
// (MTPC/MTPC.cpp:44), which fixes the order inside runs of equal Morton codes.
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <numeric>
#include <random>
#include <array>
#include "sceneManagement.h"
#include "BVH.h"
#include "pathTracing.h"

bool render_scene(std::string path, std::string filename, int N_ray_per_pixel);

using clk = std::chrono::steady_clock;
static double ms_since(clk::time_point t0) { return std::chrono::duration<double, std::milli>(clk::now() - t0).count(); }

static void wr(FILE* f, const void* p, size_t n) { if (fwrite(p, 1, n, f) != n) { perror("fwrite"); exit(2); } }

typedef std::array<double, 9> Key9;
static Key9 key_of(Face& f) { return {f.v1.x, f.v1.y, f.v1.z, f.v2.x, f.v2.y, f.v2.z, f.v3.x, f.v3.y, f.v3.z}; }

// Loads + stable-sorts; perm[k] = OBJ-order index of sorted face k.
static void load_sorted(scene_data& scene, const std::string& base, std::vector<unsigned>& codes, std::vector<int>& perm) {
    scene.read_scene(base);
    size_t n = scene.f.size();
    codes.resize(n);
    for (size_t i = 0; i < n; i++) codes[i] = scene.f[i].morton_code;
    perm.resize(n);
    std::iota(perm.begin(), perm.end(), 0);
    std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return codes[a] < codes[b]; });
    std::stable_sort(scene.f.begin(), scene.f.end(), compare);
}

static int cmd_morton(int argc, char** argv) {
    if (argc < 4) return 1;
    scene_data scene; std::vector<unsigned> codes; std::vector<int> perm;
    load_sorted(scene, argv[2], codes, perm);
    FILE* f = fopen(argv[3], "wb");
    int32_t n = (int32_t)codes.size();
    wr(f, &n, 4); wr(f, codes.data(), 4 * (size_t)n); wr(f, perm.data(), 4 * (size_t)n);
    // also the per-face centroid as narrowed to float at the getMortonCode call (sceneManagement.cpp:176-179)
    std::vector<int> inv(n); for (int k = 0; k < n; k++) inv[perm[k]] = k;
    for (int i = 0; i < n; i++) {
        Face& fc = scene.f[inv[i]];
        Vertex c = (fc.v1 + fc.v2 + fc.v3) / 3;
        float c3[3] = {(float)c.x, (float)c.y, (float)c.z};
        wr(f, c3, 12);
    }
    fclose(f);
    return 0;
}

// Time only sort + BVH ctor (MTPC.cpp:44-46), parsing excluded.
static int cmd_buildtime(int argc, char** argv) {
    if (argc < 3) return 1;
    int reps = argc > 3 ? atoi(argv[3]) : 5;
    scene_data scene; scene.read_scene(argv[2]);
    std::vector<Face> orig = scene.f;
    double best = 1e30;
    for (int r = 0; r < reps; r++) {
        scene.f = orig;
        auto t0 = clk::now();
        std::sort(scene.f.begin(), scene.f.end(), compare);
        BVH* b = new BVH(scene);
        double ms = ms_since(t0);
        delete b;
        if (ms < best) best = ms;
    }
    fprintf(stderr, "REF_BUILD_MS %.4f faces %zu\n", best, scene.f.size());
    return 0;
}

static int cmd_render(int argc, char** argv) {
    if (argc < 5) return 1;
    auto t0 = clk::now();
    render_scene(argv[2], argv[3], atoi(argv[4]));
    fprintf(stderr, "REF_RENDER_MS %.3f\n", ms_since(t0));
    return 0;
}

// image <base> <spp> <w> <h> <out>: generateImg with optional resolution override, raw doubles out.
static int cmd_image(int argc, char** argv) {
    if (argc < 7) return 1;
    scene_data scene; std::vector<unsigned> codes; std::vector<int> perm;
    load_sorted(scene, argv[2], codes, perm);
    int spp = atoi(argv[3]), w = atoi(argv[4]), h = atoi(argv[5]);
    if (w > 0) scene.camera.width = w;
    if (h > 0) scene.camera.height = h;
    BVH bvh(scene);
    image img(scene.camera.width, scene.camera.height);
    auto t0 = clk::now();
    generateImg(scene, bvh, img, spp);
    double ms = ms_since(t0);
    FILE* f = fopen(argv[6], "wb");
    int32_t hdr[3] = {scene.camera.width, scene.camera.height, spp};
    wr(f, hdr, 12);
    wr(f, img.img, sizeof(double) * 3 * (size_t)hdr[0] * hdr[1]);
    fclose(f);
    fprintf(stderr, "REF_IMAGE_MS %.3f w %d h %d spp %d\n", ms, hdr[0], hdr[1], spp);
    return 0;
}

// trace <base> <rays_in> <out>: rays_in = int32 n, then n x 6 doubles (origin, direction).
// out = int32 n, then per ray {int32 hit, int32 face (OBJ order), double t, p[3], pn[3]} packed.
static int cmd_trace(int argc, char** argv) {
    if (argc < 5) return 1;
    scene_data scene; std::vector<unsigned> codes; std::vector<int> perm;
    load_sorted(scene, argv[2], codes, perm);
    BVH bvh(scene);
    // identical geometry -> lowest sorted position (the DFS tie winner, pathTracing.cpp:355-361)
    std::map<Key9, int> first_sorted;
    for (int k = (int)scene.f.size() - 1; k >= 0; k--) first_sorted[key_of(scene.f[k])] = k;
    FILE* fi = fopen(argv[3], "rb");
    if (!fi) { perror(argv[3]); return 2; }
    int32_t n = 0;
    if (fread(&n, 4, 1, fi) != 1) return 2;
    std::vector<double> rays(6 * (size_t)n);
    if (fread(rays.data(), 8, rays.size(), fi) != rays.size()) return 2;
    fclose(fi);
    FILE* fo = fopen(argv[4], "wb");
    wr(fo, &n, 4);
    auto t0 = clk::now();
    for (int i = 0; i < n; i++) {
        Ray r;
        r.start = Vertex(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
        r.direction = Vertex(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
        intersection ret;
        bool hit = ray_intersect(r, scene, bvh, ret);
        int32_t rec[2] = {hit ? 1 : 0, -1};
        double d[7] = {0, 0, 0, 0, 0, 0, 0};
        if (hit) {
            rec[1] = perm[first_sorted[key_of(ret.f)]];
            d[0] = ret.t; d[1] = ret.p.x; d[2] = ret.p.y; d[3] = ret.p.z;
            d[4] = ret.pn.x; d[5] = ret.pn.y; d[6] = ret.pn.z;
        }
        wr(fo, rec, 8); wr(fo, d, 56);
    }
    fprintf(stderr, "REF_TRACE_MS %.3f rays %d\n", ms_since(t0), n);
    fclose(fo);
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: %s render <dir/> <name> <spp> | morton <base> <out> | buildtime <base> [reps] |\n"
                        "       image <base> <spp> <w> <h> <out> | trace <base> <rays_in> <out>\n", argv[0]);
        return 1;
    }
    std::string c = argv[1];
    if (c == "render") return cmd_render(argc, argv);
    if (c == "morton") return cmd_morton(argc, argv);
    if (c == "buildtime") return cmd_buildtime(argc, argv);
    if (c == "image") return cmd_image(argc, argv);
    if (c == "trace") return cmd_trace(argc, argv);
    return 1;
}
