// microbench.cu — the measured denominators of the roofline lines bench.py prints ("the measured L2 bandwidth
// for CB/VM", HBM for the 5 M-triangle scene): a read-only sweep with the access shape of the traversal kernels' loads (16-B
// vectors, L1 bypassed) over a working set the caller sizes — below the 126 MB L2 it measures L2 -> SM bandwidth, far above it
// HBM read bandwidth.  Not on the product path.
#include <string>
#include "context.h"

void mcpt_set_error(const char* msg);

namespace mcpt {
__global__ void __launch_bounds__(256) read_sweep_kernel(const uint4* __restrict__ buf, const size_t n_vec, const int repeats,
                                                         unsigned* __restrict__ sink) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    uint4 acc = make_uint4(0u, 0u, 0u, 0u);
    for (int r = 0; r < repeats; r++) {
        // a different rotation every repeat, so no thread re-reads the line it still holds in flight
        size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x + (size_t)r * 1315423911ull) % stride;
        for (; i < n_vec; i += stride) {
            const uint4 v = __ldcg(&buf[i]);
            acc.x ^= v.x; acc.y += v.y; acc.z ^= v.z; acc.w += v.w;
        }
    }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x9E3779B9u) *sink = acc.x;   // practically never: keeps the loads alive
}
}  // namespace mcpt

extern "C" MCPT_API int mcpt_measure_read_bandwidth(mcpt_ctx* c, int64_t working_set_bytes, int repeats, float* gbs_out) {
    if (!c || !gbs_out || working_set_bytes < 4096 || repeats < 1) { mcpt_set_error("mcpt_measure_read_bandwidth: bad argument"); return MCPT_ERR_ARG; }
    auto fail = [&](cudaError_t e, const char* what) {
        mcpt_set_error((std::string(what) + ": " + cudaGetErrorName(e)).c_str());
        return (int)MCPT_ERR_CUDA;
    };
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(e, "cudaSetDevice");
    const size_t n_vec = (size_t)working_set_bytes / 16;
    uint4* buf = nullptr; unsigned* sink = nullptr;
    if ((e = cudaMalloc((void**)&buf, n_vec * 16 + 16)) != cudaSuccess) return fail(e, "cudaMalloc");
    sink = (unsigned*)(buf + n_vec);
    cudaMemsetAsync(buf, 0x5A, n_vec * 16 + 16, c->stream);
    const int blocks = 148 * 8;
    mcpt::read_sweep_kernel<<<blocks, 256, 0, c->stream>>>(buf, n_vec, 2, sink);                 // warm: brings the set into L2
    cudaEventRecord(c->ev0, c->stream);
    mcpt::read_sweep_kernel<<<blocks, 256, 0, c->stream>>>(buf, n_vec, repeats, sink);
    cudaEventRecord(c->ev1, c->stream);
    e = cudaEventSynchronize(c->ev1);
    float ms = 0.f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    cudaFree(buf);
    if (e != cudaSuccess) return fail(e, "read_sweep_kernel");
    *gbs_out = (float)((double)n_vec * 16.0 * repeats / (ms * 1e-3) / 1e9);
    return MCPT_OK;
}
