// multigpu.cu — SURVEY.md 8(e) behind the C ABI: the SPP split over R GPUs and its ONE exchange.
//
// The reference's only parallelism is an 8-thread OpenMP loop over the samples of one pixel (MTPC/pathTracing.cpp:300-321).
// Here the samples are always cut into MCPT_NUM_CHUNKS chunks in sample order; GPU r of R in {1,2,4,8} integrates chunks
// [r*C/R, (r+1)*C/R) (scene and BVH replicated, Philox keyed by (pixel, sample)) and leaves their fixed pairwise-tree SUM in one
// FP32 buffer.  The R buffers travel to rank 0 in one grouped ncclSend/ncclRecv — straight into one preallocated R x count
// buffer, no staging copy — and K12 sums them by the same pairwise tree: the association of every float addition is the same
// for every R, so the 1-, 2-, 4- and 8-GPU images are bit-identical.  ncclReduce is NOT used: its summation order is unspecified.
//
// Two launch shapes share the code:
//   one process per GPU (torchrun, MPI):  mcpt_comm_unique_id / mcpt_comm_init / mcpt_render_dist   (ncclCommInitRank)
//   one process, all GPUs (the C++ host): mcpt_multi_*                                            (ncclCommInitAll)
// NCCL is bound at run time (dlopen of libnccl.so.2 — the copy already in the process if there is one, e.g. PyTorch's): a
// single-GPU user of libmcpt.so needs no NCCL at all, and the library never pulls a second NCCL into a torch process.
#include <dlfcn.h>
#include <nccl.h>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "context.h"

void mcpt_set_error(const char* msg);

namespace {
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);       // already in the process (PyTorch bundles one)?
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!h) return nullptr;
    bool ok = true;
    auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) ok = false; return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.Send = (decltype(api.Send))sym("ncclSend");
    api.Recv = (decltype(api.Recv))sym("ncclRecv");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    if (!ok) { dlclose(h); return nullptr; }
    api.handle = h;
    return &api;
}
int fail(int code, const std::string& m) { mcpt_set_error(m.c_str()); return code; }
int fail_nccl(NcclApi* n, ncclResult_t r, const char* what) {
    return fail(MCPT_ERR_CUDA, std::string(what) + ": NCCL " + (n && n->GetErrorString ? n->GetErrorString(r) : "error"));
}
int fail_cuda(cudaError_t e, const char* what) { return fail(MCPT_ERR_CUDA, std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")"); }
#define NC(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) return fail_nccl(N, r_, #call); } while (0)
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(e_, #call); } while (0)
bool pow2_upto8(int n) { return n == 1 || n == 2 || n == 4 || n == 8; }

cudaError_t grow_f(float*& p, size_t& cap, size_t bytes) {
    if (bytes <= cap && p) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc((void**)&p, bytes ? bytes : 1);
    if (e == cudaSuccess) cap = bytes;
    return e;
}
void merge_stats(mcpt_stats& into, const mcpt_stats& s, bool first) {
    if (first) { into = s; return; }
    into.samples += s.samples; into.rays += s.rays; into.node_visits += s.node_visits; into.tri_tests += s.tri_tests;
    into.kernel_launches += s.kernel_launches; into.trace_launches += s.trace_launches;
    if (s.render_ms > into.render_ms) into.render_ms = s.render_ms;          // ranks run side by side: the slowest one counts
    if (s.integrate_ms > into.integrate_ms) into.integrate_ms = s.integrate_ms;
    if (s.trace_ms > into.trace_ms) into.trace_ms = s.trace_ms;
    if (s.build_ms > into.build_ms) into.build_ms = s.build_ms;
}
}  // namespace

struct mcpt_comm { ncclComm_t comm = nullptr; int world = 1, rank = 0; };

void mcpt_comm_release(mcpt_ctx* c) {
    if (!c || !c->comm) return;
    NcclApi* N = nccl_api();
    if (N && c->comm->comm) N->CommDestroy(c->comm->comm);
    delete c->comm;
    c->comm = nullptr;
}

extern "C" {

MCPT_API int mcpt_comm_unique_id(char id_out[MCPT_COMM_ID_BYTES]) {
    static_assert(sizeof(ncclUniqueId) <= MCPT_COMM_ID_BYTES, "ncclUniqueId grew");
    if (!id_out) return fail(MCPT_ERR_ARG, "mcpt_comm_unique_id: null argument");
    NcclApi* N = nccl_api();
    if (!N) return fail(MCPT_ERR_CUDA, "mcpt_comm_unique_id: libnccl.so.2 not found (needed only for more than one GPU)");
    ncclUniqueId id;
    NC(N->GetUniqueId(&id));
    std::memset(id_out, 0, MCPT_COMM_ID_BYTES);
    std::memcpy(id_out, &id, sizeof(id));
    return MCPT_OK;
}

MCPT_API int mcpt_comm_init(mcpt_ctx* c, int world, int rank, const char id[MCPT_COMM_ID_BYTES]) {
    if (!c || !id) return fail(MCPT_ERR_ARG, "mcpt_comm_init: null argument");
    if (!pow2_upto8(world) || rank < 0 || rank >= world)
        return fail(MCPT_ERR_ARG, "mcpt_comm_init: world must be 1, 2, 4 or 8 (it divides the sample chunks) and 0 <= rank < world");
    mcpt_comm_release(c);
    mcpt_comm* m = new mcpt_comm;
    m->world = world; m->rank = rank;
    if (world > 1) {
        NcclApi* N = nccl_api();
        if (!N) { delete m; return fail(MCPT_ERR_CUDA, "mcpt_comm_init: libnccl.so.2 not found"); }
        CU(cudaSetDevice(c->device));
        ncclUniqueId uid;
        std::memcpy(&uid, id, sizeof(uid));
        ncclResult_t r = N->CommInitRank(&m->comm, world, uid, rank);
        if (r != ncclSuccess) { delete m; return fail_nccl(N, r, "ncclCommInitRank"); }
    }
    c->comm = m;
    return MCPT_OK;
}

MCPT_API int mcpt_comm_destroy(mcpt_ctx* c) {
    if (!c) return fail(MCPT_ERR_ARG, "mcpt_comm_destroy: null ctx");
    mcpt_comm_release(c);
    return MCPT_OK;
}

// The exchange + fixed-order sum shared by both launch shapes.  comms/ctxs of the ranks THIS process drives (one for
// mcpt_render_dist, all of them for mcpt_multi_render); partial[i] = rank i's tree sum (rank 0's already sits in stacked[0]).
static int exchange_and_sum(NcclApi* N, std::vector<mcpt_ctx*>& ctxs, std::vector<ncclComm_t>& comms, const std::vector<int>& ranks,
                            int world, const std::vector<float*>& partial, size_t count, int spp, float* exchange_ms) {
    mcpt_ctx* root = nullptr;
    for (size_t i = 0; i < ctxs.size(); i++) if (ranks[i] == 0) root = ctxs[i];
    if (root) { CU(cudaSetDevice(root->device)); CU(cudaEventRecord(root->ev0, root->stream)); }
    if (world > 1) {
        NC(N->GroupStart());
        for (size_t i = 0; i < ctxs.size(); i++) {
            if (ranks[i] == 0) {
                for (int r = 1; r < world; r++) NC(N->Recv(root->d_stacked + (size_t)r * count, count, ncclFloat, r, comms[i], root->stream));
            } else {
                NC(N->Send(partial[i], count, ncclFloat, 0, comms[i], ctxs[i]->stream));
            }
        }
        NC(N->GroupEnd());
    }
    if (root) {
        CU(cudaSetDevice(root->device));
        CU(cudaEventRecord(root->ev1, root->stream));
        CU(grow_f(root->d_image, root->image_cap, sizeof(float) * count));
        const int rc = mcpt_sum_partials(root, root->d_stacked, world, (int64_t)count, 1.0f / (float)spp, root->d_image);
        if (rc) return rc;
        CU(cudaStreamSynchronize(root->stream));
        if (exchange_ms) CU(cudaEventElapsedTime(exchange_ms, root->ev0, root->ev1));
    }
    for (size_t i = 0; i < ctxs.size(); i++)
        if (ranks[i] != 0) { CU(cudaSetDevice(ctxs[i]->device)); CU(cudaStreamSynchronize(ctxs[i]->stream)); }   // the send buffer is reused next call
    return MCPT_OK;
}

MCPT_API int mcpt_render_dist(mcpt_ctx* c, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags, float* rgb_out,
                              mcpt_stats* st) {
    if (!c || !cam) return fail(MCPT_ERR_ARG, "mcpt_render_dist: null argument");
    if (!c->comm) return fail(MCPT_ERR_ARG, "mcpt_render_dist: call mcpt_comm_init first");
    if (width <= 0) width = cam->width;
    if (height <= 0) height = cam->height;
    if (width <= 0 || height <= 0 || spp <= 0) return fail(MCPT_ERR_ARG, "mcpt_render_dist: width, height and spp must be positive");
    const int world = c->comm->world, rank = c->comm->rank, per = MCPT_NUM_CHUNKS / world;
    const size_t count = (size_t)width * height * 3;
    NcclApi* N = world > 1 ? nccl_api() : nullptr;
    CU(cudaSetDevice(c->device));
    CU(grow_f(c->d_stacked, c->stacked_cap, sizeof(float) * count * (rank == 0 ? (size_t)world : 1)));
    mcpt_stats s{};
    int rc = mcpt_render_chunks(c, cam, width, height, spp, rank * per, (rank + 1) * per, mode, flags, c->d_stacked, &s);
    if (rc) return rc;
    std::vector<mcpt_ctx*> ctxs{c}; std::vector<ncclComm_t> comms{c->comm->comm}; std::vector<int> ranks{rank}; std::vector<float*> partial{c->d_stacked};
    float ex_ms = 0.f;
    rc = exchange_and_sum(N, ctxs, comms, ranks, world, partial, count, spp, &ex_ms);
    if (rc) return rc;
    if (rank == 0 && rgb_out) {
        CU(cudaMemcpyAsync(rgb_out, c->d_image, sizeof(float) * count, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
    }
    s.exchange_ms = ex_ms; s.kernel_launches += rank == 0 ? 1 : 0;
    if (st) *st = s;
    return MCPT_OK;
}

MCPT_API const float* mcpt_device_image(mcpt_ctx* c) { return c ? c->d_image : nullptr; }

}  // extern "C"

// ---- one process, all GPUs ----------------------------------------------------------------------------------------
struct mcpt_multi {
    std::vector<mcpt_ctx*> ctx;
    std::vector<ncclComm_t> comm;
    std::vector<int> ranks;
    std::vector<float*> partial;   // per rank: where its tree sum goes (rank 0: its slot of the stacked buffer)
};

namespace {
// run fn(i) for every context on its own host thread (every C-ABI call blocks on its own stream); first error wins
template <class F> int for_each_ctx(mcpt_multi* m, F fn) {
    const int n = (int)m->ctx.size();
    std::vector<int> rc(n, 0);
    std::vector<std::string> msg(n);
    auto body = [&](int i) { rc[i] = fn(i); if (rc[i]) msg[i] = mcpt_last_error(); };   // mcpt_last_error is thread-local
    std::vector<std::thread> pool;
    for (int i = 1; i < n; i++) pool.emplace_back(body, i);
    body(0);
    for (auto& t : pool) t.join();
    for (int i = 0; i < n; i++) if (rc[i]) { mcpt_set_error(("GPU " + std::to_string(i) + ": " + msg[i]).c_str()); return rc[i]; }
    return MCPT_OK;
}
}  // namespace

extern "C" {

MCPT_API int mcpt_multi_create(const int* devices, int n_dev, mcpt_multi** out) {
    if (!out) return fail(MCPT_ERR_ARG, "mcpt_multi_create: null out");
    if (!pow2_upto8(n_dev)) return fail(MCPT_ERR_ARG, "mcpt_multi_create: the number of GPUs must be 1, 2, 4 or 8 (it divides the sample chunks)");
    mcpt_multi* m = new mcpt_multi;
    std::vector<int> devs(n_dev);
    for (int i = 0; i < n_dev; i++) devs[i] = devices ? devices[i] : i;
    for (int i = 0; i < n_dev; i++) {
        mcpt_ctx* c = nullptr;
        const int rc = mcpt_create(devs[i], &c);
        if (rc) { for (mcpt_ctx* p : m->ctx) mcpt_destroy(p); delete m; return rc; }
        m->ctx.push_back(c); m->ranks.push_back(i);
    }
    m->comm.assign(n_dev, nullptr);
    m->partial.assign(n_dev, nullptr);
    if (n_dev > 1) {
        NcclApi* N = nccl_api();
        ncclResult_t r = N ? N->CommInitAll(m->comm.data(), n_dev, devs.data()) : ncclSystemError;
        if (!N || r != ncclSuccess) {
            for (mcpt_ctx* p : m->ctx) mcpt_destroy(p);
            delete m;
            return N ? fail_nccl(N, r, "ncclCommInitAll") : fail(MCPT_ERR_CUDA, "mcpt_multi_create: libnccl.so.2 not found");
        }
    }
    *out = m;
    return MCPT_OK;
}

MCPT_API void mcpt_multi_destroy(mcpt_multi* m) {
    if (!m) return;
    NcclApi* N = nccl_api();
    for (size_t i = 0; i < m->ctx.size(); i++) {
        if (m->comm[i] && N) { cudaSetDevice(m->ctx[i]->device); cudaStreamSynchronize(m->ctx[i]->stream); N->CommDestroy(m->comm[i]); }
    }
    for (mcpt_ctx* c : m->ctx) mcpt_destroy(c);
    delete m;
}

MCPT_API int mcpt_multi_size(const mcpt_multi* m) { return m ? (int)m->ctx.size() : 0; }
MCPT_API mcpt_ctx* mcpt_multi_ctx(mcpt_multi* m, int i) { return (m && i >= 0 && i < (int)m->ctx.size()) ? m->ctx[i] : nullptr; }

MCPT_API int mcpt_multi_set_seed(mcpt_multi* m, uint64_t seed) {
    if (!m) return fail(MCPT_ERR_ARG, "mcpt_multi_set_seed: null argument");
    for (mcpt_ctx* c : m->ctx) mcpt_set_seed(c, seed);
    return MCPT_OK;
}

MCPT_API int mcpt_multi_upload_scene(mcpt_multi* m, const mcpt_scene_desc* scene) {
    if (!m || !scene) return fail(MCPT_ERR_ARG, "mcpt_multi_upload_scene: null argument");
    return for_each_ctx(m, [&](int i) { return mcpt_upload_scene(m->ctx[i], scene); });
}

MCPT_API int mcpt_multi_load_scene_gpu(mcpt_multi* m, const char* path, const char* name, mcpt_camera* camera_out) {
    if (!m || !path || !name) return fail(MCPT_ERR_ARG, "mcpt_multi_load_scene_gpu: null argument");
    std::vector<mcpt_camera> cams(m->ctx.size());
    const int rc = for_each_ctx(m, [&](int i) { return mcpt_scene_load_gpu(m->ctx[i], path, name, &cams[i], nullptr); });
    if (!rc && camera_out) *camera_out = cams[0];
    return rc;
}

MCPT_API int mcpt_multi_build_bvh(mcpt_multi* m, float* ms_out) {
    if (!m) return fail(MCPT_ERR_ARG, "mcpt_multi_build_bvh: null argument");
    std::vector<float> ms(m->ctx.size(), 0.f);
    const int rc = for_each_ctx(m, [&](int i) { return mcpt_build_bvh(m->ctx[i], &ms[i]); });
    if (!rc && ms_out) { *ms_out = 0.f; for (float v : ms) if (v > *ms_out) *ms_out = v; }
    return rc;
}

// Leaves the mean image in GPU 0's d_image.
static int multi_render_device(mcpt_multi* m, const mcpt_camera* cam, int& width, int& height, int spp, int mode, int flags, mcpt_stats* st) {
    if (!m || !cam) return fail(MCPT_ERR_ARG, "mcpt_multi_render: null argument");
    if (width <= 0) width = cam->width;
    if (height <= 0) height = cam->height;
    if (width <= 0 || height <= 0 || spp <= 0) return fail(MCPT_ERR_ARG, "mcpt_multi_render: width, height and spp must be positive");
    const int world = (int)m->ctx.size(), per = MCPT_NUM_CHUNKS / world;
    const size_t count = (size_t)width * height * 3;
    for (int i = 0; i < world; i++) {
        mcpt_ctx* c = m->ctx[i];
        CU(cudaSetDevice(c->device));
        CU(grow_f(c->d_stacked, c->stacked_cap, sizeof(float) * count * (i == 0 ? (size_t)world : 1)));
        m->partial[i] = c->d_stacked;
    }
    std::vector<mcpt_stats> stats(world);
    int rc = for_each_ctx(m, [&](int i) {
        return mcpt_render_chunks(m->ctx[i], cam, width, height, spp, i * per, (i + 1) * per, mode, flags, m->partial[i], &stats[i]);
    });
    if (rc) return rc;
    float ex_ms = 0.f;
    rc = exchange_and_sum(world > 1 ? nccl_api() : nullptr, m->ctx, m->comm, m->ranks, world, m->partial, count, spp, &ex_ms);
    if (rc) return rc;
    if (st) {
        for (int i = 0; i < world; i++) merge_stats(*st, stats[i], i == 0);
        st->exchange_ms = ex_ms; st->kernel_launches += 1;
    }
    return MCPT_OK;
}

MCPT_API int mcpt_multi_render(mcpt_multi* m, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags, float* rgb_out,
                               mcpt_stats* st) {
    if (!rgb_out) return fail(MCPT_ERR_ARG, "mcpt_multi_render: null output");
    const int rc = multi_render_device(m, cam, width, height, spp, mode, flags, st);
    if (rc) return rc;
    mcpt_ctx* c = m->ctx[0];
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(rgb_out, c->d_image, sizeof(float) * (size_t)width * height * 3, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return MCPT_OK;
}

MCPT_API int mcpt_multi_render_png(mcpt_multi* m, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags,
                                   const char* file, mcpt_stats* st) {
    if (!file) return fail(MCPT_ERR_ARG, "mcpt_multi_render_png: null file name");
    int rc = multi_render_device(m, cam, width, height, spp, mode, flags, st);
    if (rc) return rc;
    mcpt_ctx* c = m->ctx[0];
    std::vector<uint8_t> png((size_t)mcpt_png_size(width, height));
    rc = mcpt_encode_png(c, c->d_image, width, height, png.data(), (int64_t)png.size());
    if (rc) return rc;
    FILE* f = std::fopen(file, "wb");
    if (!f) return fail(MCPT_ERR_IO, std::string("cannot open ") + file);
    bool ok = std::fwrite(png.data(), 1, png.size(), f) == png.size();
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) return fail(MCPT_ERR_IO, "short write");
    return MCPT_OK;
}

}  // extern "C"
