// TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// Driver/exporter linked against the UNMODIFIED reference objects (compiled by oracle/build_ref.sh
// from /root/reference/MTPC/*.cpp where they lie).  It only calls the reference's own functions:
//   render_scene            MTPC/MTPC.cpp:35
//   scene_data::read_scene  MTPC/sceneManagement.cpp:264
//   compare                 MTPC/sceneManagement.cpp:311
//   BVH::BVH                MTPC/BVH.cpp:37
//   ray_intersect           MTPC/pathTracing.cpp:382
//   generateImg             MTPC/pathTracing.cpp:274
//   imshow (+ svpng)        MTPC/MTPC.cpp:10, MTPC/svpng.inc:77
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
void imshow(double* SRC, int WIDTH, int HEIGHT, std::string filename, int N_ray_per_pixel);   // MTPC/MTPC.cpp:10

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
    // optional: the reference's own output stage on the same image -> <png_prefix>-SPP<spp>.png (never fclose'd by imshow;
    // the normal process exit below flushes it)
    if (argc > 7) imshow(img.img, scene.camera.width, scene.camera.height, argv[7], spp);
    return 0;
}

// png <w> <h> <doubles_in> <png_prefix> <spp>: imshow + svpng on a caller-supplied double image (w*h*3 raw doubles).
static int cmd_png(int argc, char** argv) {
    if (argc < 7) return 1;
    const int w = atoi(argv[2]), h = atoi(argv[3]);
    std::vector<double> img(3 * (size_t)w * h);
    FILE* fi = fopen(argv[4], "rb");
    if (!fi) { perror(argv[4]); return 2; }
    if (fread(img.data(), 8, img.size(), fi) != img.size()) return 2;
    fclose(fi);
    imshow(img.data(), w, h, argv[5], atoi(argv[6]));
    return 0;
}

// scene <base> <out>: what read_scene parsed, per face in OBJ order: v1 v2 v3 (9 doubles), vn1 vn2 vn3 (9), vt1 vt2 vt3 (6),
// then the material name as int32 length + bytes.  Header: int32 n.  (scene.f is NOT sorted here.)
static int cmd_scene(int argc, char** argv) {
    if (argc < 4) return 1;
    scene_data scene;
    scene.read_scene(argv[2]);
    FILE* f = fopen(argv[3], "wb");
    int32_t n = (int32_t)scene.f.size();
    wr(f, &n, 4);
    for (int i = 0; i < n; i++) {
        Face& fc = scene.f[i];
        double d[24] = {fc.v1.x, fc.v1.y, fc.v1.z, fc.v2.x, fc.v2.y, fc.v2.z, fc.v3.x, fc.v3.y, fc.v3.z,
                        fc.vn1.x, fc.vn1.y, fc.vn1.z, fc.vn2.x, fc.vn2.y, fc.vn2.z, fc.vn3.x, fc.vn3.y, fc.vn3.z,
                        fc.vt1.vtx, fc.vt1.vty, fc.vt2.vtx, fc.vt2.vty, fc.vt3.vtx, fc.vt3.vty};
        wr(f, d, sizeof(d));
        int32_t len = (int32_t)fc.material.size();
        wr(f, &len, 4); wr(f, fc.material.data(), (size_t)len);
    }
    fclose(f);
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
                        "       image <base> <spp> <w> <h> <out> [png_prefix] | trace <base> <rays_in> <out> |\n"
                        "       png <w> <h> <doubles_in> <png_prefix> <spp> | scene <base> <out>\n", argv[0]);
        return 1;
    }
    std::string c = argv[1];
    if (c == "render") return cmd_render(argc, argv);
    if (c == "morton") return cmd_morton(argc, argv);
    if (c == "buildtime") return cmd_buildtime(argc, argv);
    if (c == "image") return cmd_image(argc, argv);
    if (c == "trace") return cmd_trace(argc, argv);
    if (c == "png") return cmd_png(argc, argv);
    if (c == "scene") return cmd_scene(argc, argv);
    return 1;
}
