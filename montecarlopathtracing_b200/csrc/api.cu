// api.cu — the C ABI of include/mcpt.h (device side): context, scene upload, seams 1-3.
// No CPU fallback: every entry point that computes needs a CUDA device with sm_100a code.
#include <algorithm>
#include <atomic>
#include <thread>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <type_traits>
#include <unordered_map>
#include <vector>
#include "context.h"

void mcpt_set_error(const char* msg);  // host/mcpt_error.cpp

using namespace mcpt;

namespace {
int fail_cuda(cudaError_t e, const char* what) {
    std::string m = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    mcpt_set_error(m.c_str());
    return MCPT_ERR_CUDA;
}
int fail_arg(const char* what) { mcpt_set_error(what); return MCPT_ERR_ARG; }
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)

template <class T> cudaError_t dev_alloc(T*& p, size_t count) {
    if (p) { cudaFree(p); p = nullptr; }
    return cudaMalloc((void**)&p, sizeof(T) * (count ? count : 1));
}
template <class T> cudaError_t grow(T*& p, size_t& cap, size_t bytes) {
    if (bytes <= cap && p) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc((void**)&p, bytes ? bytes : 1);
    if (e == cudaSuccess) cap = bytes;
    return e;
}
void free_scene(mcpt_ctx* c) {   // forgets the scene; the arenas stay (released by mcpt_destroy)
    c->sb = SceneBuffers();
    c->d_materials = nullptr; c->d_textures = nullptr; c->d_tex_bytes = nullptr; c->d_lights = nullptr;
    c->d_light_tris = nullptr; c->d_light_cdf = nullptr;
    c->scene_loaded = false;
}
// Face::calAera, MTPC/sceneManagement.cpp:399-406 (law of cosines), in double like the reference
double face_area(const double* p) {
    auto len = [](double x, double y, double z) { return std::sqrt(x * x + y * y + z * z); };
    const double a = len(p[3] - p[0], p[4] - p[1], p[5] - p[2]), b = len(p[6] - p[0], p[7] - p[1], p[8] - p[2]),
                 c = len(p[6] - p[3], p[7] - p[4], p[8] - p[5]);
    const double cos_c = (a * a + b * b - c * c) / (2 * a * b);
    return a * b * std::sqrt(1 - cos_c * cos_c) / 2;
}
DeviceScene device_scene(const mcpt_ctx* c) {
    DeviceScene d{};
    d.n_faces = c->sb.n_faces; d.n_materials = c->n_materials; d.n_textures = c->n_textures; d.n_lights = c->n_lights;
    d.root = MCPT_WIDE ? 0 : c->sb.root; d.qnodes4 = c->sb.qnodes4; d.area0 = c->area0;
    d.n_wide = c->sb.n_wide; d.wide_levels = c->sb.wide_levels;
    d.nodes = c->sb.nodes; d.qnodes = c->sb.qnodes; d.qgrid = c->sb.qgrid; d.tri = c->sb.tri; d.shade = c->sb.shade; d.leaf_face = c->sb.leaf_face; d.pos = c->sb.pos;
    d.materials = c->d_materials; d.textures = c->d_textures; d.tex_bytes = c->d_tex_bytes;
    d.lights = c->d_lights; d.light_tris = c->d_light_tris; d.light_cdf = c->d_light_cdf;
    return d;
}

int check_tables(const mcpt_scene_desc* d, const char* who) {
    auto bad = [&](const char* what) { return fail_arg((std::string(who) + ": " + what).c_str()); };
    if (d->n_materials <= 0 || !d->materials) return bad("no materials");
    for (int l = 0; l < d->n_lights; l++)
        if (d->lights[l].mat_id < 0 || d->lights[l].mat_id >= d->n_materials) return bad("light mat_id out of range");
    for (int m = 0; m < d->n_materials; m++)
        if (d->materials[m].tex_id >= d->n_textures) return bad("tex_id out of range");
    for (int i = 0; i < d->n_textures; i++)
        if (d->textures[i].height <= 0 || d->textures[i].width <= 0 || !d->textures[i].bgr) return bad("bad texture");
    return MCPT_OK;
}

// every per-face device array of a scene of n faces (geometry, build scratch, trees)
int alloc_scene(mcpt_ctx* c, int n) {
    SceneBuffers& s = c->sb;
    s.n_faces = n;
    const size_t N = (size_t)n;
    s.ploc_temp_bytes = ploc_scratch_bytes(n);
    s.cub_temp_bytes = lbvh_sort_temp_bytes(n);
    { const char* v = getenv("MCPT_BVH"); s.use_ploc = !(v && std::strcmp(v, "lbvh") == 0); }
    { const char* v = getenv("MCPT_LEAF_PAIRS"); s.fold_pairs = !(v && std::strcmp(v, "0") == 0); }
    // one allocation, carved (a scene of n faces needs ~30 arrays; 30 cudaMalloc/cudaFree pairs cost more than a small build)
    char* base = nullptr;
    size_t used = 0;
    auto carve = [&](auto*& p, size_t count) {
        using T = std::remove_reference_t<decltype(*p)>;
        if (base) p = (T*)(base + used);
        used += (sizeof(T) * (count ? count : 1) + 255) & ~(size_t)255;
    };
    auto carve_bytes = [&](void*& p, size_t bytes) {
        if (base) p = base + used;
        used += ((bytes ? bytes : 1) + 255) & ~(size_t)255;
    };
    for (int pass = 0; pass < 2; pass++) {
        used = 0;
        carve(s.pos, 9 * N); carve(s.nrm, 9 * N); carve(s.uv, 6 * N); carve(s.mat_id, N);
        carve(s.codes_obj, N); carve(s.codes_sorted, N); carve(s.iota, N); carve(s.leaf_face, N);
        carve(s.leaf_parent, N); carve(s.flags, N); carve(s.nodes, N); carve(s.tri, N);
        carve(s.shade, N); carve(s.root_box, 6); carve(s.qnodes, N); carve(s.qgrid, 1);
        carve(s.qnodes4, N); carve(s.w4_queue, N); carve(s.w4_ctr, 16); carve(s.w4_cand, n - 1 <= W4_SMALL_MAX_N ? N : 1);
        carve(s.nodes2, N); carve(s.cl_lo[0], N); carve(s.cl_lo[1], N); carve(s.cl_hi[0], N); carve(s.cl_hi[1], N);
        carve(s.ploc_flag, N);
        carve_bytes(s.ploc_temp, s.ploc_temp_bytes); carve_bytes(s.ploc_state, 64); carve_bytes(s.cub_temp, s.cub_temp_bytes);
        if (pass == 0) { CU(c->scene_arena.reserve(used)); base = c->scene_arena.base; s.arena = base; }
    }
    return MCPT_OK;
}

// materials, textures and lights; lpos/lnrm[l] = the faces of light l's material in OBJ order (9 doubles each)
int upload_tables(mcpt_ctx* c, const mcpt_scene_desc* d, const std::vector<std::vector<double>>& lpos,
                  const std::vector<std::vector<double>>& lnrm) {
    cudaStream_t st = c->stream;
    // materials: light_id = light_map lookup (the last "mtlname" line naming the material wins, sceneManagement.cpp:256)
    std::vector<DMaterial> mats(d->n_materials);
    for (int m = 0; m < d->n_materials; m++) {
        const mcpt_material& src = d->materials[m];
        DMaterial& t = mats[m];
        for (int k = 0; k < 3; k++) { t.kd[k] = (float)src.kd[k]; t.ks[k] = (float)src.ks[k]; }
        t.kd_norm = (float)std::sqrt(src.kd[0] * src.kd[0] + src.kd[1] * src.kd[1] + src.kd[2] * src.kd[2]);
        t.ks_norm = (float)std::sqrt(src.ks[0] * src.ks[0] + src.ks[1] * src.ks[1] + src.ks[2] * src.ks[2]);
        t.Ns = (float)src.Ns; t.Ni = (float)src.Ni; t.tex_id = src.tex_id; t.light_id = -1;
    }
    for (int l = 0; l < d->n_lights; l++) mats[d->lights[l].mat_id].light_id = l;
    // textures, packed back to back
    std::vector<DTexture> texs(d->n_textures);
    std::vector<unsigned char> tex_bytes;
    for (int i = 0; i < d->n_textures; i++) {
        const mcpt_texture& t = d->textures[i];
        texs[i].height = t.height; texs[i].width = t.width; texs[i].offset = tex_bytes.size();
        tex_bytes.insert(tex_bytes.end(), t.bgr, t.bgr + (size_t)t.height * t.width * 3);
    }
    // lights: area CDF over the light material's faces in OBJ order (MTPC/pathTracing.cpp:177-184), built once
    std::vector<DLight> lights(d->n_lights);
    std::vector<LightTri> ltris;
    std::vector<float> cdf;
    std::vector<double> areas(d->n_lights, 0.0);
    for (int l = 0; l < d->n_lights; l++) {
        DLight& L = lights[l];
        L.mat_id = d->lights[l].mat_id; L.first = (int)ltris.size(); L.count = 0;
        for (int k = 0; k < 3; k++) L.radiance[k] = (float)d->lights[l].radiance[k];
        double total = 0;
        float blo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, bhi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
        for (size_t j = 0; j < lpos[l].size() / 9; j++) {
            const double* p = &lpos[l][9 * j]; const double* q = &lnrm[l][9 * j];
            for (int k = 0; k < 9; k++) { blo[k % 3] = std::min(blo[k % 3], (float)p[k]); bhi[k % 3] = std::max(bhi[k % 3], (float)p[k]); }
            total += face_area(p);
            cdf.push_back((float)total);
            LightTri t;
            t.p0n0x = make_float4((float)p[0], (float)p[1], (float)p[2], (float)q[0]);
            t.p1n0y = make_float4((float)p[3], (float)p[4], (float)p[5], (float)q[1]);
            t.p2n0z = make_float4((float)p[6], (float)p[7], (float)p[8], (float)q[2]);
            t.n1n2x = make_float4((float)q[3], (float)q[4], (float)q[5], (float)q[6]);
            t.n2yz = make_float4((float)q[7], (float)q[8], 0.f, 0.f);
            ltris.push_back(t);
            L.count++;
        }
        areas[l] = total; L.area = (float)total;
        // box of the light material's triangles, pushed out past the rounding of the traversal triangles (v0 + e1 in FP32)
        for (int k = 0; k < 3; k++) {
            const float pad = 1.0e-5f * std::max(std::fabs(blo[k]), std::fabs(bhi[k])) + 1.0e-6f;
            L.box_lo[k] = blo[k] - pad; L.box_hi[k] = bhi[k] + pad;
        }
    }
    c->area0 = d->n_lights > 0 ? (float)areas[0] : 0.f;
    for (int l = 0; l < d->n_lights; l++) {   // entries reachable by rnd <= area0 (quirk Q6): bounds the binary search
        DLight& L = lights[l];
        int m = 0;
        while (m < L.count && !(cdf[L.first + m] > c->area0)) m++;
        L.n_search = m < L.count ? m + 1 : L.count;
    }
    c->n_materials = d->n_materials; c->n_textures = d->n_textures; c->n_lights = d->n_lights;
    {
        char* base = nullptr;
        size_t used = 0;
        auto carve = [&](auto*& p, size_t count) {
            using T = std::remove_reference_t<decltype(*p)>;
            if (base) p = (T*)(base + used);
            used += (sizeof(T) * (count ? count : 1) + 255) & ~(size_t)255;
        };
        for (int pass = 0; pass < 2; pass++) {
            used = 0;
            carve(c->d_materials, mats.size()); carve(c->d_textures, texs.size()); carve(c->d_tex_bytes, tex_bytes.size());
            carve(c->d_lights, lights.size()); carve(c->d_light_tris, ltris.size()); carve(c->d_light_cdf, cdf.size());
            if (pass == 0) { CU(c->tables_arena.reserve(used)); base = c->tables_arena.base; }
        }
    }
    CU(cudaMemcpyAsync(c->d_materials, mats.data(), sizeof(DMaterial) * mats.size(), cudaMemcpyHostToDevice, st));
    if (!texs.empty()) CU(cudaMemcpyAsync(c->d_textures, texs.data(), sizeof(DTexture) * texs.size(), cudaMemcpyHostToDevice, st));
    if (!tex_bytes.empty()) CU(cudaMemcpyAsync(c->d_tex_bytes, tex_bytes.data(), tex_bytes.size(), cudaMemcpyHostToDevice, st));
    if (!lights.empty()) CU(cudaMemcpyAsync(c->d_lights, lights.data(), sizeof(DLight) * lights.size(), cudaMemcpyHostToDevice, st));
    if (!ltris.empty()) CU(cudaMemcpyAsync(c->d_light_tris, ltris.data(), sizeof(LightTri) * ltris.size(), cudaMemcpyHostToDevice, st));
    if (!cdf.empty()) CU(cudaMemcpyAsync(c->d_light_cdf, cdf.data(), sizeof(float) * cdf.size(), cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));  // the staging vectors die here
    c->scene_loaded = true;
    return MCPT_OK;
}

// mcpt_ingest_obj after the text is on its way to c->d_text (ev0 recorded): parse, flatten, tables
int ingest_resident(mcpt_ctx* c, int64_t n_bytes, const mcpt_scene_desc* d, const char* const* names, float* ms_out) {
    cudaStream_t st = c->stream;
    ObjIngest g;
    g.d_text = c->d_text; g.n_bytes = (unsigned)n_bytes;
    g.arena_text = &c->ingest_text; g.arena_stmt = &c->ingest_stmt;
    const bool timing = getenv("MCPT_INGEST_TIMING") != nullptr;   // host wall clock per phase, to stderr
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[ingest] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    int e = obj_ingest_parse(g, st);
    lap("H2D wait + count/scan/parse");
    if (e) return fail_cuda((cudaError_t)e, "obj_ingest_parse");
    auto obj_error = [&](unsigned long long w) {
        static const char* what[] = {"", "only triangles are supported (as in the reference)", "face uses unknown material",
                                     "face index out of range", "number not readable the way atof reads it (hex / inf / nan / > 40 digits)"};
        const unsigned code = (unsigned)(w & 0xFF);
        std::string m = std::string("obj: ") + what[code < 5 ? code : 0] + " (line at byte offset " + std::to_string(w >> 8) + ")";
        mcpt_set_error(m.c_str());
        return (int)MCPT_ERR_IO;
    };
    if (g.n_f == 0) {                                   // errors wait for stage 2: an unknown material on an earlier line comes first
        if (g.h_status[0] != ~0ull) return obj_error(g.h_status[0]);
        mcpt_set_error("obj: no faces");
        return MCPT_ERR_IO;
    }
    std::vector<std::string> um_names;
    e = obj_ingest_names(g, um_names, st);
    if (e) return fail_cuda((cudaError_t)e, "obj_ingest_names");
    std::unordered_map<std::string, int> by_name;
    for (int m = 0; m < d->n_materials; m++)
        if (names[m]) by_name[names[m]] = m;              // a repeated newmtl name keeps one entry (scene_loader.cpp)
    std::vector<int> um_mat(um_names.size(), -1);
    for (size_t i = 0; i < um_names.size(); i++) {
        const auto it = by_name.find(um_names[i]);
        if (it != by_name.end()) um_mat[i] = it->second;
    }
    lap("usemtl names");
    if (int rc = alloc_scene(c, (int)g.n_f)) return rc;
    lap("alloc_scene");
    SceneBuffers& s = c->sb;
    e = obj_ingest_flatten(g, um_mat.data(), s.pos, s.nrm, s.uv, s.mat_id, st);
    if (e) return fail_cuda((cudaError_t)e, "obj_ingest_flatten");
    if (g.h_status[0] != ~0ull) { free_scene(c); return obj_error(g.h_status[0]); }
    lap("flatten");
    // light faces: material ids back to the host (4 B per face), index lists per light, one gather per light
    const int n = (int)g.n_f;
    std::vector<int> mat((size_t)n);
    CU(cudaMemcpyAsync(mat.data(), s.mat_id, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    std::vector<std::vector<double>> lpos(d->n_lights), lnrm(d->n_lights);
    for (int l = 0; l < d->n_lights; l++) {
        std::vector<int> idx;
        for (int i = 0; i < n; i++) if (mat[i] == d->lights[l].mat_id) idx.push_back(i);
        if (idx.empty()) continue;
        // the parsed statements are dead after the flatten: their arena doubles as the gather's scratch
        const size_t out_off = (4 * idx.size() + 255) & ~(size_t)255;
        CU(c->ingest_stmt.reserve(out_off + 144 * idx.size()));
        int* d_idx = (int*)c->ingest_stmt.base; double* d_out = (double*)(c->ingest_stmt.base + out_off);
        CU(cudaMemcpyAsync(d_idx, idx.data(), 4 * idx.size(), cudaMemcpyHostToDevice, st));
        e = launch_gather_faces(d_idx, (int)idx.size(), s.pos, s.nrm, d_out, st);
        std::vector<double> packed(18 * idx.size());
        if (!e) e = (int)cudaMemcpyAsync(packed.data(), d_out, 8 * packed.size(), cudaMemcpyDeviceToHost, st);
        if (!e) e = (int)cudaStreamSynchronize(st);
        if (e) return fail_cuda((cudaError_t)e, "gather light faces");
        lpos[l].resize(9 * idx.size()); lnrm[l].resize(9 * idx.size());
        for (size_t j = 0; j < idx.size(); j++) {
            std::memcpy(&lpos[l][9 * j], &packed[18 * j], 72);
            std::memcpy(&lnrm[l][9 * j], &packed[18 * j + 9], 72);
        }
    }
    lap("light faces");
    if (int rc = upload_tables(c, d, lpos, lnrm)) return rc;
    lap("tables");
    CU(cudaEventRecord(c->ev1, st));
    CU(cudaEventSynchronize(c->ev1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->last_ingest_launches = g.launches;
    if (ms_out) *ms_out = ms;
    return MCPT_OK;
}

}  // namespace

extern "C" {

MCPT_API int mcpt_create(int device, mcpt_ctx** out) {
    if (!out) return fail_arg("mcpt_create: null out");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDeviceCount (no CUDA device: this library has no CPU fallback)");
    if (device < 0 || device >= n) return fail_arg("mcpt_create: device index out of range");
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        mcpt_set_error("mcpt_create: kernels are built for sm_100a (B200) only");
        return MCPT_ERR_CUDA;
    }
    mcpt_ctx* c = new mcpt_ctx;
    c->device = device;
    CU(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
    c->stream = c->own_stream;
    CU(cudaEventCreate(&c->ev0)); CU(cudaEventCreate(&c->ev1)); CU(cudaEventCreate(&c->ev2)); CU(cudaEventCreate(&c->ev3));
    CU(dev_alloc(c->d_counters, 8));
    CU(cudaStreamCreateWithFlags(&c->build_streams.side, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&c->build_streams.ev_fork, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->build_streams.ev_join, cudaEventDisableTiming));
    *out = c;
    return MCPT_OK;
}

MCPT_API void mcpt_destroy(mcpt_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    mcpt_comm_release(c);
    cudaFree(c->d_stacked);
    free_scene(c);
    cudaFree(c->d_counters); cudaFree(c->d_planes); cudaFree(c->d_image); cudaFree(c->d_io); cudaFree(c->d_dirs); cudaFree(c->d_wf); cudaFree(c->d_png); cudaFree(c->d_text); c->ingest_text.release(); c->ingest_stmt.release(); c->scene_arena.release(); c->tables_arena.release(); c->wfh.destroy();
    for (int k = 0; k < PIN_RING; k++) { if (c->h_pin[k]) cudaFreeHost(c->h_pin[k]); if (c->pin_ev[k]) cudaEventDestroy(c->pin_ev[k]); }
    cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1); cudaEventDestroy(c->ev2); cudaEventDestroy(c->ev3);
    cudaStreamDestroy(c->own_stream);
    if (c->build_streams.side) cudaStreamDestroy(c->build_streams.side);
    if (c->build_streams.ev_fork) cudaEventDestroy(c->build_streams.ev_fork);
    if (c->build_streams.ev_join) cudaEventDestroy(c->build_streams.ev_join);
    delete c;
}

MCPT_API int mcpt_set_stream(mcpt_ctx* c, void* s) {
    if (!c) return fail_arg("mcpt_set_stream: null ctx");
    c->stream = (cudaStream_t)s;   // NULL is a stream too: CUDA's legacy default stream (what torch.cuda.current_stream() is by default)
    return MCPT_OK;
}
MCPT_API int mcpt_use_own_stream(mcpt_ctx* c) {
    if (!c) return fail_arg("mcpt_use_own_stream: null ctx");
    c->stream = c->own_stream;
    return MCPT_OK;
}
MCPT_API int mcpt_set_seed(mcpt_ctx* c, uint64_t seed) {
    if (!c) return fail_arg("mcpt_set_seed: null ctx");
    c->seed = seed;
    return MCPT_OK;
}

MCPT_API int mcpt_upload_scene(mcpt_ctx* c, const mcpt_scene_desc* d) {
    if (!c || !d) return fail_arg("mcpt_upload_scene: null argument");
    if (d->n_faces <= 0 || !d->pos || !d->nrm || !d->uv || !d->mat_id) return fail_arg("mcpt_upload_scene: empty scene");
    if (int rc = check_tables(d, "mcpt_upload_scene")) return rc;
    const int n = d->n_faces;
    for (int i = 0; i < n; i++)
        if (d->mat_id[i] < 0 || d->mat_id[i] >= d->n_materials) return fail_arg("mcpt_upload_scene: mat_id out of range");
    CU(cudaSetDevice(c->device));
    free_scene(c);
    if (int rc = alloc_scene(c, n)) return rc;
    SceneBuffers& s = c->sb;
    const size_t N = (size_t)n;
    cudaStream_t st = c->stream;
    CU(cudaMemcpyAsync(s.pos, d->pos, 72 * N, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.nrm, d->nrm, 72 * N, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.uv, d->uv, 48 * N, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.mat_id, d->mat_id, 4 * N, cudaMemcpyHostToDevice, st));
    // faces of every light's material, OBJ order
    std::vector<std::vector<double>> lpos(d->n_lights), lnrm(d->n_lights);
    for (int l = 0; l < d->n_lights; l++)
        for (int i = 0; i < n; i++) {
            if (d->mat_id[i] != d->lights[l].mat_id) continue;
            lpos[l].insert(lpos[l].end(), d->pos + 9 * (size_t)i, d->pos + 9 * (size_t)i + 9);
            lnrm[l].insert(lnrm[l].end(), d->nrm + 9 * (size_t)i, d->nrm + 9 * (size_t)i + 9);
        }
    return upload_tables(c, d, lpos, lnrm);
}

/* Scene ingest on the GPU: see include/mcpt.h and csrc/obj_ingest.cu. */
MCPT_API int mcpt_ingest_obj(mcpt_ctx* c, const char* text, int64_t n_bytes, const mcpt_scene_desc* d, const char* const* names,
                             float* ms_out) {
    if (!c || !text || !d || !names) return fail_arg("mcpt_ingest_obj: null argument");
    if (n_bytes <= 0 || n_bytes >= (int64_t)0xFFFFFF00ll) return fail_arg("mcpt_ingest_obj: text must be 1 byte .. 4 GiB");
    if (int rc = check_tables(d, "mcpt_ingest_obj")) return rc;
    CU(cudaSetDevice(c->device));
    free_scene(c);
    cudaStream_t st = c->stream;
    CU(cudaEventRecord(c->ev0, st));
    CU(grow(c->d_text, c->text_cap, (size_t)n_bytes + 16));   // the staging loop reads whole words
    CU(cudaMemcpyAsync(c->d_text, text, (size_t)n_bytes, cudaMemcpyHostToDevice, st));
    return ingest_resident(c, n_bytes, d, names, ms_out);
}

/* The same from a file: 8 MiB blocks read by 4 threads into a pinned ring, each block's H2D copy overlapping the reads. */
MCPT_API int mcpt_ingest_obj_file(mcpt_ctx* c, const char* file, const mcpt_scene_desc* d, const char* const* names, float* ms_out) {
    if (!c || !file || !d || !names) return fail_arg("mcpt_ingest_obj_file: null argument");
    if (int rc = check_tables(d, "mcpt_ingest_obj_file")) return rc;
    const int fd = ::open(file, O_RDONLY);
    if (fd < 0) { mcpt_set_error((std::string("cannot open ") + file).c_str()); return MCPT_ERR_IO; }
    struct Closer { int fd; ~Closer() { ::close(fd); } } closer{fd};
    struct stat sb;
    if (::fstat(fd, &sb) != 0) { mcpt_set_error((std::string("cannot stat ") + file).c_str()); return MCPT_ERR_IO; }
    const long long n_bytes = (long long)sb.st_size;
    if (n_bytes <= 0) { mcpt_set_error((std::string(file) + ": no faces").c_str()); return MCPT_ERR_IO; }
    if (n_bytes >= 0xFFFFFF00ll) return fail_arg("mcpt_ingest_obj_file: file must be < 4 GiB");
    const bool timing = getenv("MCPT_INGEST_TIMING") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[ingest] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    CU(cudaSetDevice(c->device));
    free_scene(c);
    lap("free_scene");
    cudaStream_t st = c->stream;
    const size_t BLOCK = 8u << 20;
    for (int k = 0; k < PIN_RING; k++) {
        if (!c->h_pin[k]) CU(cudaHostAlloc((void**)&c->h_pin[k], BLOCK, cudaHostAllocDefault));
        if (!c->pin_ev[k]) CU(cudaEventCreateWithFlags(&c->pin_ev[k], cudaEventDisableTiming));
    }
    lap("pinned ring");
    CU(cudaEventRecord(c->ev0, st));
    CU(grow(c->d_text, c->text_cap, (size_t)n_bytes + 16));   // the staging loop reads whole words
    // PIN_RING readers, reader k owning pinned buffer k and blocks k, k + PIN_RING, ...: pread (page cache -> pinned, one
    // core copies ~10 GB/s) runs in parallel and overlaps the H2D copies, which all go to the context's stream
    const size_t n_blocks = ((size_t)n_bytes + BLOCK - 1) / BLOCK;
    std::atomic<int> failed{0};
    auto reader = [&](int k) {
        if (cudaSetDevice(c->device) != cudaSuccess) { failed = 2; return; }
        for (size_t blk = (size_t)k; blk < n_blocks && !failed; blk += PIN_RING) {
            const size_t off = blk * BLOCK, len = std::min(BLOCK, (size_t)n_bytes - off);
            if (cudaEventSynchronize(c->pin_ev[k]) != cudaSuccess) { failed = 2; return; }   // the copy that last used this buffer
            size_t got = 0;
            while (got < len) {
                const ssize_t r = ::pread(fd, c->h_pin[k] + got, len - got, (off_t)(off + got));
                if (r <= 0) { failed = 1; return; }
                got += (size_t)r;
            }
            if (cudaMemcpyAsync(c->d_text + off, c->h_pin[k], len, cudaMemcpyHostToDevice, st) != cudaSuccess ||
                cudaEventRecord(c->pin_ev[k], st) != cudaSuccess) { failed = 2; return; }
        }
    };
    {
        std::vector<std::thread> pool;
        const int n_readers = (int)std::min<size_t>(PIN_RING, n_blocks);
        for (int k = 1; k < n_readers; k++) pool.emplace_back(reader, k);
        reader(0);
        for (auto& t : pool) t.join();
    }
    if (failed == 1) { mcpt_set_error((std::string("cannot read ") + file).c_str()); return MCPT_ERR_IO; }
    if (failed) return fail_cuda(cudaGetLastError(), "ingest file copy");
    lap("file read + H2D enqueue");
    return ingest_resident(c, n_bytes, d, names, ms_out);
}

MCPT_API int64_t mcpt_scene_face_count(mcpt_ctx* c) {
    if (!c || !c->scene_loaded) { fail_arg("mcpt_scene_face_count: no scene"); return MCPT_ERR_ARG; }
    return c->sb.n_faces;
}

MCPT_API int mcpt_get_scene_arrays(mcpt_ctx* c, double* pos, double* nrm, double* uv, int32_t* mat_id) {
    if (!c || !c->scene_loaded) return fail_arg("mcpt_get_scene_arrays: no scene");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    const size_t N = (size_t)c->sb.n_faces;
    if (pos) CU(cudaMemcpy(pos, c->sb.pos, 72 * N, cudaMemcpyDeviceToHost));
    if (nrm) CU(cudaMemcpy(nrm, c->sb.nrm, 72 * N, cudaMemcpyDeviceToHost));
    if (uv) CU(cudaMemcpy(uv, c->sb.uv, 48 * N, cudaMemcpyDeviceToHost));
    if (mat_id) CU(cudaMemcpy(mat_id, c->sb.mat_id, 4 * N, cudaMemcpyDeviceToHost));
    return MCPT_OK;
}

MCPT_API int mcpt_build_bvh(mcpt_ctx* c, float* ms_out) {
    if (!c || !c->scene_loaded) return fail_arg("mcpt_build_bvh: upload a scene first");
    CU(cudaSetDevice(c->device));
    c->sb.built = false;
    c->sb.w4_resident_blocks = c->coop_blocks[0]; c->sb.ploc_resident_blocks = c->coop_blocks[1];   // occupancy queries, once per context
    CU(cudaEventRecord(c->ev0, c->stream));
    int launches = 0;
    const int e = lbvh_build(c->sb, c->stream, c->build_streams, &launches);
    if (e) return fail_cuda((cudaError_t)e, "lbvh_build");
    c->coop_blocks[0] = c->sb.w4_resident_blocks; c->coop_blocks[1] = c->sb.ploc_resident_blocks;
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaEventSynchronize(c->ev1));
    // post-build verdicts (device flags): bad vertices; quality pass stalled; wide tree deeper than the traversal stack serves
    unsigned w4[8] = {};
    CU(cudaMemcpy(w4, c->sb.w4_ctr, sizeof(w4), cudaMemcpyDeviceToHost));
    if (w4[5]) { c->sb.built = false; return fail_arg("mcpt_build_bvh: a vertex coordinate is not finite or beyond +-1e18"); }
    bool fallback = false;
    if (!c->sb.walk_is_radix) {
        cudaError_t pe = cudaSuccess;
        if (!ploc_ok(c->sb, &pe)) { if (pe != cudaSuccess) return fail_cuda(pe, "ploc state"); fallback = true; }
    }
#if MCPT_WIDE
    auto wide_bad = [&]() { return w4[3] || !w4[2]; };
    auto too_deep = [&]() { return w4[4] > MCPT_MAX_WIDE_LEVELS; };
    if (!fallback && !c->sb.walk_is_radix && (wide_bad() || too_deep())) fallback = true;
#endif
    if (fallback) {   // walk the radix tree instead (its depth is bounded by the key length)
        const int fe = lbvh_fallback_to_radix(c->sb, c->stream, &launches);
        if (fe) return fail_cuda((cudaError_t)fe, "lbvh_fallback_to_radix");
        CU(cudaEventRecord(c->ev1, c->stream));
        CU(cudaEventSynchronize(c->ev1));
        CU(cudaMemcpy(w4, c->sb.w4_ctr, sizeof(w4), cudaMemcpyDeviceToHost));
    }
    CU(cudaEventElapsedTime(&c->last_build_ms, c->ev0, c->ev1));
    if (ms_out) *ms_out = c->last_build_ms;
    c->last_build_launches = launches;
    c->sb.n_wide = (int)w4[0]; c->sb.wide_levels = (int)w4[4];
#if MCPT_WIDE
    if (wide_bad()) { c->sb.built = false; return fail_cuda(cudaErrorLaunchTimeout, "wide_build: the breadth-first collapse did not finish"); }
    if (too_deep()) {
        c->sb.built = false;
        return fail_arg("mcpt_build_bvh: the tree is deeper than the traversal stack serves (more than 21 wide levels): degenerate geometry");
    }
#endif
    return MCPT_OK;
}

MCPT_API int mcpt_get_morton(mcpt_ctx* c, uint32_t* codes_obj, uint32_t* codes_sorted, int32_t* face_sorted) {
    if (!c || !c->sb.built) return fail_arg("mcpt_get_morton: build the BVH first");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    const size_t b = 4 * (size_t)c->sb.n_faces;
    if (codes_obj) CU(cudaMemcpy(codes_obj, c->sb.codes_obj, b, cudaMemcpyDeviceToHost));
    if (codes_sorted) CU(cudaMemcpy(codes_sorted, c->sb.codes_sorted, b, cudaMemcpyDeviceToHost));
    if (face_sorted) CU(cudaMemcpy(face_sorted, c->sb.leaf_face, b, cudaMemcpyDeviceToHost));
    return MCPT_OK;
}

MCPT_API int mcpt_get_nodes(mcpt_ctx* c, float* nodes16, int32_t* root) {
    if (!c || !c->sb.built) return fail_arg("mcpt_get_nodes: build the BVH first");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    const int n_nodes = c->sb.n_faces > 1 ? c->sb.n_faces - 1 : 1;
    if (nodes16) CU(cudaMemcpy(nodes16, c->sb.nodes, sizeof(BvhNode) * (size_t)n_nodes, cudaMemcpyDeviceToHost));
    if (root) *root = 0;
    return MCPT_OK;
}

MCPT_API int mcpt_get_wide_tree(mcpt_ctx* c, uint32_t* nodes16, int32_t* n_wide, float* grid8) {
    if (!c || !c->sb.built) return fail_arg("mcpt_get_wide_tree: build the BVH first");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    unsigned w4[4] = {0, 0, 0, 0};
    CU(cudaMemcpy(w4, c->sb.w4_ctr, sizeof(w4), cudaMemcpyDeviceToHost));
    if (n_wide) *n_wide = (int32_t)w4[0];
    const int cap = c->sb.n_faces > 1 ? c->sb.n_faces - 1 : 1;
    if (nodes16) CU(cudaMemcpy(nodes16, c->sb.qnodes4, sizeof(QNode4) * (size_t)cap, cudaMemcpyDeviceToHost));
    if (grid8) CU(cudaMemcpy(grid8, c->sb.qgrid, sizeof(QGrid), cudaMemcpyDeviceToHost));
    return MCPT_OK;
}

MCPT_API int mcpt_get_traversal_tree(mcpt_ctx* c, float* nodes16, int32_t* root) {
    if (!c || !c->sb.built) return fail_arg("mcpt_get_traversal_tree: build the BVH first");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    const int n_nodes = c->sb.n_faces > 1 ? c->sb.n_faces - 1 : 1;
    const BvhNode* src = c->sb.walk_is_radix ? c->sb.nodes : c->sb.nodes2;
    if (nodes16) CU(cudaMemcpy(nodes16, src, sizeof(BvhNode) * (size_t)n_nodes, cudaMemcpyDeviceToHost));
    if (root) *root = c->sb.root;
    return MCPT_OK;
}

namespace {
void fill_stats(mcpt_ctx* c, mcpt_stats* st, double samples, const unsigned long long cnt[3], float ms, int launches) {
    if (!st) return;
    st->samples = samples; st->rays = (double)cnt[0]; st->node_visits = (double)cnt[1]; st->tri_tests = (double)cnt[2];
    st->render_ms = ms; st->build_ms = c->last_build_ms; st->build_launches = c->last_build_launches; st->exchange_ms = 0.f; st->kernel_launches = launches; st->integrate_ms = 0.f;
    st->trace_ms = 0.f; st->trace_launches = 0;
}
// generateImg's camera basis (MTPC/pathTracing.cpp:276-294), in double on the host
DCamera make_camera(const mcpt_camera* cam, int width, int height) {
    DCamera d{};
    const double PI = 3.1415926;  // MTPC/pathTracing.h:11
    double up[3] = {cam->up[0], cam->up[1], cam->up[2]};
    const double ul = std::sqrt(up[0] * up[0] + up[1] * up[1] + up[2] * up[2]);
    for (double& v : up) v /= ul;
    double dir[3];
    for (int k = 0; k < 3; k++) dir[k] = cam->lookat[k] - cam->eye[k];
    const double l = std::sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
    const double dy = std::tan(cam->fovy / 2 / 180 * PI) * l;
    const double dx = dy / height * width;
    const double pdx = 2 * dx / width, pdy = 2 * dy / height;
    double X[3] = {dir[1] * up[2] - up[1] * dir[2], up[0] * dir[2] - dir[0] * up[2], dir[0] * up[1] - up[0] * dir[1]};
    const double xl = std::sqrt(X[0] * X[0] + X[1] * X[1] + X[2] * X[2]);
    for (double& v : X) v /= xl;
    for (int k = 0; k < 3; k++) {
        d.eye[k] = cam->eye[k];
        d.start[k] = cam->lookat[k] - X[k] * dx + up[k] * dy;
        d.pdx[k] = X[k] * pdx;
        d.pdy[k] = up[k] * pdy;
    }
    d.width = width; d.height = height;
    return d;
}
// Enqueue the integrator for chunks [cb, ce) into c->d_planes (ce-cb planes). No sync.
int enqueue_chunks(mcpt_ctx* c, const mcpt_camera* cam, int width, int height, int spp, int cb, int ce, int mode, int flags,
                   int* launches) {
    if (mode != MCPT_MODE_MEGAKERNEL && mode != MCPT_MODE_WAVEFRONT) return fail_arg("render: unknown mode");
    if (getenv("MCPT_NO_SHADOW_EARLY_OUT")) flags |= MCPT_FLAG_NO_SHADOW_EARLY_OUT;   // developer A/B switch (same image either way)
    if (flags & MCPT_FLAG_JITTER) flags |= MCPT_FLAG_NO_PRIMARY_REUSE;                  // every sample has its own primary ray
    if (flags & MCPT_FLAG_MIS_BALANCE) flags |= MCPT_FLAG_ANYHIT_SHADOW | MCPT_FLAG_NO_PRIMARY_REUSE;
    const int nch = ce - cb;
    const size_t plane = (size_t)width * height * 3;
    CU(grow(c->d_planes, c->planes_cap, sizeof(float) * plane * nch));
    const size_t dirs_bytes = (sizeof(float) * plane + 255) & ~(size_t)255;   // directions, then K7a's FP64 positions
    CU(grow(c->d_dirs, c->dirs_cap, dirs_bytes + sizeof(double) * plane));
    c->wfh.iterations = 0;
    RenderParams P{};
    P.scene = device_scene(c);
    P.cam = make_camera(cam, width, height);
    P.cam_dirs = c->d_dirs;
    const int er = launch_raygen(P.cam, c->d_dirs, (double*)((char*)c->d_dirs + dirs_bytes), c->stream);
    if (er) return fail_cuda((cudaError_t)er, "raygen kernels launch");
    if (launches) *launches += 2;
    P.spp = spp; P.chunk_begin = cb; P.n_chunks = nch;
    // enough threads to fill 148 SMs several times over, otherwise as many samples per thread as possible
    // Degree of parallelism = pixels x (chunks / chunks per thread).  Megakernel: enough threads to fill 148 SMs several
    // times over.  Wavefront: enough path slots that one wf_trace launch carries ~2 M rays per pipe (a slot emits one
    // extension ray and one shadow ray per light and iteration): with fewer, a launch is mostly its drain phase — the
    // queue is empty and every warp waits for its longest ray — and cornell-box at 1024^2 ran 28 % slower.
    int g = nch;
    long long want = 600000LL;
    if (mode == MCPT_MODE_WAVEFRONT) {
        const char* v = getenv("MCPT_WF_SLOTS");
        want = v && *v ? atoll(v) : std::max(600000LL, 6000000LL / (c->n_lights + 1));
        if (want > (1LL << 25)) want = 1LL << 25;
    }
    while (g > 1 && (long long)width * height * (nch / g) < want) g >>= 1;
    if (mode == MCPT_MODE_WAVEFRONT) {
        // at least four slots per pixel where that fits (2^25 slots): wf_shade hands a pixel's chunks out dynamically, which only
        // shortens the render's tail when a pixel has more than one slot to share them (3840 x 2160 and veach-mis had one)
        const char* v = getenv("MCPT_WF_MIN_GROUPS");
        int min_groups = v && *v ? atoi(v) : 4;   // measured 1 / 2 / 4: 3840x2160 3135 / 3277 / 3296 Msamples/s, veach-mis 712 / 729 / 733, 5 M triangles 946 / 945 / 967
        const char* m = getenv("MCPT_WF_MAX_SLOTS");
        const long long max_slots = m && *m ? atoll(m) : (1LL << 25);
        while (g > 1 && nch / g < min_groups && (long long)width * height * (nch / (g >> 1)) <= max_slots) g >>= 1;
    }
    P.chunks_per_thread = g;
    P.flags = flags;
    P.div_tiles_x = make_fastdiv((unsigned)((width + 7) >> 3));
    P.div_tiles = make_fastdiv((unsigned)(((width + 7) >> 3) * ((height + 3) >> 2)));
    P.seed_lo = (unsigned)c->seed; P.seed_hi = (unsigned)(c->seed >> 32);
    P.planes = c->d_planes; P.counters = c->d_counters;
    CU(cudaMemsetAsync(c->d_counters, 0, 8 * sizeof(unsigned long long), c->stream));
    CU(cudaEventRecord(c->ev3, c->stream));
    if (mode == MCPT_MODE_WAVEFRONT) {
        const int n_slots = wavefront_slots(P);
        if (n_slots < 0) return fail_arg("render: image too large for one wavefront pass (more than 2^26 path slots)");
        if (c->n_lights > 32) return fail_arg("render: wavefront mode supports at most 32 lights");
        CU(grow(c->d_wf, c->wf_cap, wavefront_scratch_bytes(n_slots, c->n_lights, (long long)width * height)));
        const int e = launch_wavefront(P, flags, c->d_wf, c->wfh, c->stream, launches);
        if (e) return fail_cuda((cudaError_t)e, "wavefront loop");
    } else {
        const int e = launch_megakernel(P, flags, c->stream);
        if (e) return fail_cuda((cudaError_t)e, "pt_megakernel launch");
        if (launches) *launches += 1;
    }
    CU(cudaEventRecord(c->ev2, c->stream));
    return MCPT_OK;
}
// after the stream is synchronised: integrator time and, in wavefront mode, the summed wf_trace time
int fill_kernel_times(mcpt_ctx* c, mcpt_stats* st) {
    if (!st) return MCPT_OK;
    CU(cudaEventElapsedTime(&st->integrate_ms, c->ev3, c->ev2));
    double sum = 0;
    for (int i = 0; i < c->wfh.iterations; i++) {
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, c->wfh.trace_events[2 * i], c->wfh.trace_events[2 * i + 1]));
        sum += ms;
    }
    st->trace_ms = (float)sum; st->trace_launches = c->wfh.iterations;
    return MCPT_OK;
}
int check_render_args(mcpt_ctx* c, const mcpt_camera* cam, int& width, int& height, int spp) {
    if (!c || !cam) return fail_arg("render: null argument");
    if (!c->sb.built) return fail_arg("render: build the BVH first");
    if (width <= 0) width = cam->width;
    if (height <= 0) height = cam->height;
    if (width <= 0 || height <= 0 || spp <= 0) return fail_arg("render: width, height and spp must be positive");
    return MCPT_OK;
}
}  // namespace

MCPT_API int mcpt_trace_device(mcpt_ctx* c, const float* d_org_dir, int n, mcpt_hit* d_hits, int flags, mcpt_stats* st) {
    if (!c || !c->sb.built) return fail_arg("mcpt_trace: build the BVH first");
    if (n < 0 || (n > 0 && (!d_org_dir || !d_hits))) return fail_arg("mcpt_trace: bad buffers");
    CU(cudaSetDevice(c->device));
    CU(cudaMemsetAsync(c->d_counters, 0, 8 * sizeof(unsigned long long), c->stream));
    CU(cudaEventRecord(c->ev0, c->stream));
    const int e = launch_trace(device_scene(c), d_org_dir, n, d_hits, (flags & MCPT_FLAG_COUNT) != 0, c->d_counters, c->stream);
    if (e) return fail_cuda((cudaError_t)e, "trace_kernel launch");
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaEventSynchronize(c->ev1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    unsigned long long cnt[4];
    CU(cudaMemcpy(cnt, c->d_counters, sizeof(cnt), cudaMemcpyDeviceToHost));
    cnt[0] = (unsigned long long)n;
    fill_stats(c, st, 0.0, cnt, ms, 1);
    return MCPT_OK;
}

MCPT_API int mcpt_trace(mcpt_ctx* c, const float* org_dir, int n, mcpt_hit* hits_out, mcpt_stats* st) {
    if (!c || !c->sb.built) return fail_arg("mcpt_trace: build the BVH first");
    if (n < 0 || (n > 0 && (!org_dir || !hits_out))) return fail_arg("mcpt_trace: bad buffers");
    if (n == 0) return MCPT_OK;
    CU(cudaSetDevice(c->device));
    const size_t in_b = 24 * (size_t)n, out_b = sizeof(mcpt_hit) * (size_t)n;
    const size_t in_pad = (in_b + 255) & ~(size_t)255;
    CU(grow(c->d_io, c->io_cap, in_pad + out_b));
    float* d_in = (float*)c->d_io;
    mcpt_hit* d_out = (mcpt_hit*)((char*)c->d_io + in_pad);
    CU(cudaMemcpyAsync(d_in, org_dir, in_b, cudaMemcpyHostToDevice, c->stream));
    const int rc = mcpt_trace_device(c, d_in, n, d_out, MCPT_FLAG_COUNT, st);
    if (rc) return rc;
    CU(cudaMemcpyAsync(hits_out, d_out, out_b, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return MCPT_OK;
}

MCPT_API int mcpt_sum_partials(mcpt_ctx* c, const float* d_partials, int n, int64_t count, float scale, float* d_out) {
    if (!c || !d_partials || !d_out || count < 0) return fail_arg("mcpt_sum_partials: bad argument");
    if (!(n == 1 || n == 2 || n == 4 || n == 8)) return fail_arg("mcpt_sum_partials: n must be 1, 2, 4 or 8");
    CU(cudaSetDevice(c->device));
    const int e = launch_sum_planes(d_partials, n, (long long)count, scale, d_out, c->stream);
    if (e) return fail_cuda((cudaError_t)e, "sum_planes launch");
    return MCPT_OK;
}

MCPT_API int mcpt_render_chunks(mcpt_ctx* c, const mcpt_camera* cam, int width, int height, int spp, int cb, int ce, int mode,
                                int flags, float* d_partial, mcpt_stats* st) {
    int rc = check_render_args(c, cam, width, height, spp);
    if (rc) return rc;
    const int nch = ce - cb;
    if (cb < 0 || ce > MCPT_NUM_CHUNKS || nch < 1 || (nch & (nch - 1)) != 0 || cb % nch != 0)
        return fail_arg("mcpt_render_chunks: [chunk_begin, chunk_end) must be an aligned power-of-two range within 0..MCPT_NUM_CHUNKS");
    if (!d_partial) return fail_arg("mcpt_render_chunks: null output");
    CU(cudaSetDevice(c->device));
    int launches = 0;
    CU(cudaEventRecord(c->ev0, c->stream));
    rc = enqueue_chunks(c, cam, width, height, spp, cb, ce, mode, flags, &launches);
    if (rc) return rc;
    const long long count = (long long)width * height * 3;
    const int e = launch_sum_planes(c->d_planes, nch, count, 1.0f, d_partial, c->stream);
    if (e) return fail_cuda((cudaError_t)e, "sum_planes launch");
    launches++;
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaEventSynchronize(c->ev1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    unsigned long long cnt[4];
    CU(cudaMemcpy(cnt, c->d_counters, sizeof(cnt), cudaMemcpyDeviceToHost));
    const long long s0 = (long long)cb * spp / MCPT_NUM_CHUNKS, s1 = (long long)ce * spp / MCPT_NUM_CHUNKS;
    fill_stats(c, st, (double)width * height * (double)(s1 - s0), cnt, ms, launches);
    return fill_kernel_times(c, st);
}

// Shared by mcpt_render and mcpt_render_png: the mean image is left in c->d_image; ev0..ev1 bracket the kernels.
static int render_device(mcpt_ctx* c, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags, int* launches) {
    const size_t count = (size_t)width * height * 3;
    CU(grow(c->d_image, c->image_cap, sizeof(float) * count));
    CU(cudaEventRecord(c->ev0, c->stream));
    const int rc = enqueue_chunks(c, cam, width, height, spp, 0, MCPT_NUM_CHUNKS, mode, flags, launches);
    if (rc) return rc;
    const int e = launch_sum_planes(c->d_planes, MCPT_NUM_CHUNKS, (long long)count, 1.0f / (float)spp, c->d_image, c->stream);
    if (e) return fail_cuda((cudaError_t)e, "sum_planes launch");
    (*launches)++;
    CU(cudaEventRecord(c->ev1, c->stream));
    return MCPT_OK;
}
static int finish_render_stats(mcpt_ctx* c, mcpt_stats* st, int width, int height, int spp, int launches) {
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    unsigned long long cnt[8];
    CU(cudaMemcpy(cnt, c->d_counters, sizeof(cnt), cudaMemcpyDeviceToHost));
    fill_stats(c, st, (double)width * height * (double)spp, cnt, ms, launches);
    if (cnt[3] && getenv("MCPT_DIAG"))   // wavefront + MCPT_FLAG_COUNT: how full the warps are when they issue the two step kinds
        std::fprintf(stderr, "[diag] node steps: %.2f lanes per warp issue (%llu issues); triangle tests: %.2f lanes per issue (%llu); "
                     "%.2f node issues and %.3f tri issues per loop iteration (%llu iterations)\n",
                     (double)cnt[1] / (double)cnt[3], cnt[3], cnt[4] ? (double)cnt[2] / (double)cnt[4] : 0.0, cnt[4],
                     (double)cnt[3] / (double)cnt[5], (double)cnt[4] / (double)cnt[5], cnt[5]),
        std::fprintf(stderr, "[diag] per node-step issue: %.2f lanes idle (no ray), %.2f finished but holding a parked leaf, %.2f at a second leaf\n",
                     (double)cnt[6] / (double)cnt[3], (double)cnt[7] / (double)cnt[3],
                     32.0 - ((double)cnt[1] + (double)cnt[6] + (double)cnt[7]) / (double)cnt[3]);
    return fill_kernel_times(c, st);
}

MCPT_API int mcpt_render(mcpt_ctx* c, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags, float* rgb_out,
                         mcpt_stats* st) {
    int rc = check_render_args(c, cam, width, height, spp);
    if (rc) return rc;
    if (!rgb_out) return fail_arg("mcpt_render: null output");
    CU(cudaSetDevice(c->device));
    int launches = 0;
    rc = render_device(c, cam, width, height, spp, mode, flags, &launches);
    if (rc) return rc;
    CU(cudaMemcpyAsync(rgb_out, c->d_image, sizeof(float) * (size_t)width * height * 3, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return finish_render_stats(c, st, width, height, spp, launches);
}

MCPT_API int64_t mcpt_png_size(int width, int height) {
    if (width <= 0 || height <= 0) return 0;
    return (int64_t)png_encoded_size(width, height);
}

MCPT_API int mcpt_encode_png(mcpt_ctx* c, const float* d_rgb, int width, int height, uint8_t* png_out, int64_t capacity) {
    if (!c || !d_rgb || !png_out || width <= 0 || height <= 0) return fail_arg("mcpt_encode_png: bad argument");
    if (width > 21844) return fail_arg("mcpt_encode_png: width > 21844 does not fit svpng's one-block-per-scanline framing");
    const size_t bytes = png_encoded_size(width, height);
    if (capacity < (int64_t)bytes) return fail_arg("mcpt_encode_png: output buffer smaller than mcpt_png_size()");
    CU(cudaSetDevice(c->device));
    CU(grow(c->d_png, c->png_cap, bytes + 64));
    const int e = launch_png_encode(d_rgb, width, height, c->d_png + 64, c->d_png, c->stream);   // first 64 B: scratch
    if (e) return fail_cuda((cudaError_t)e, "png encode launch");
    CU(cudaMemcpyAsync(png_out, c->d_png + 64, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return MCPT_OK;
}

MCPT_API int mcpt_render_png(mcpt_ctx* c, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags, const char* file,
                             mcpt_stats* st) {
    int rc = check_render_args(c, cam, width, height, spp);
    if (rc) return rc;
    if (!file) return fail_arg("mcpt_render_png: null file name");
    CU(cudaSetDevice(c->device));
    int launches = 0;
    rc = render_device(c, cam, width, height, spp, mode, flags, &launches);
    if (rc) return rc;
    std::vector<uint8_t> png((size_t)mcpt_png_size(width, height));
    rc = mcpt_encode_png(c, c->d_image, width, height, png.data(), (int64_t)png.size());   // synchronises the stream
    if (rc) return rc;
    launches += 4;
    rc = finish_render_stats(c, st, width, height, spp, launches);
    if (rc) return rc;
    FILE* f = std::fopen(file, "wb");
    if (!f) { mcpt_set_error((std::string("cannot open ") + file).c_str()); return MCPT_ERR_IO; }
    bool ok = std::fwrite(png.data(), 1, png.size(), f) == png.size();
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) { mcpt_set_error("short write"); return MCPT_ERR_IO; }
    return MCPT_OK;
}

}  // extern "C"
