// ploc.cu — quality pass of seam 1: re-cluster the Morton-ordered leaves into the traversal tree (K4a).
//
// The Karras radix tree (lbvh.cu K2/K3) splits wherever the Morton codes differ; on cornell-box a ray visits 30 of its
// nodes where a surface-area-guided tree needs 12 (measured: DESIGN.md "tree quality"), and traversal — not the build —
// is 99.9 % of a render.  This pass builds the tree that traversal actually walks by parallel locally-ordered
// clustering (Meister & Bittner 2018) over the SAME Morton-sorted leaf sequence the radix tree is built on:
//   repeat until one cluster is left:
//     nn      every cluster looks at its RADIUS neighbours on either side in the (Morton-ordered) cluster array and picks
//             the one whose merged box has the smallest surface area (ties: lowest index, which guarantees a mutual pair);
//     flags   mutually-nearest pairs (i < j) become a new internal node at i, j dies;
//     scan    one exclusive scan of packed (survivors, merges) gives every survivor its new position (order preserved)
//             and every merge its node index (so the tree does not depend on scheduling);
//     scatter writes the new nodes (64-B BvhNode: both child boxes + child references) and the compacted cluster array.
// Once <= 1024 clusters remain a single CTA finishes the job in shared memory.  Node boxes are exact unions of the
// padded FP32 leaf boxes of K3, so no refit is needed.  Leaf order, leaf indices and the Morton codes are untouched.
#include <cub/block/block_scan.cuh>
#include <cub/device/device_scan.cuh>
#include "lbvh.h"

namespace mcpt {

#ifndef PLOC_RADIUS
#define PLOC_RADIUS 8
#endif
#define PLOC_TPB 256
#define PLOC_FINISH 1024

struct PlocState { int m; int node_base; };   // clusters alive, internal nodes created so far

__device__ __forceinline__ float merged_area(const float4 alo, const float4 ahi, const float4 blo, const float4 bhi) {
    const float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y),
                dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dy), __fmul_rn(dy, dz)), __fmul_rn(dz, dx));   // symmetric in (a, b), no FMA
}

// lo.w carries the cluster's reference (internal node index >= 0, leaf ~k < 0) as int bits.
__device__ __forceinline__ void write_node(BvhNode* nodes, int id, const float4 alo, const float4 ahi, const float4 blo, const float4 bhi) {
    BvhNode nd;
    nd.q[0] = make_float4(alo.x, ahi.x, alo.y, ahi.y);
    nd.q[1] = make_float4(blo.x, bhi.x, blo.y, bhi.y);
    nd.q[2] = make_float4(alo.z, ahi.z, blo.z, bhi.z);
    nd.q[3] = make_float4(alo.w, blo.w, __int_as_float(-1), 0.0f);
    nodes[id] = nd;
}

// ---- global-memory iteration -------------------------------------------------------------------------------
__global__ void __launch_bounds__(PLOC_TPB) ploc_nn_kernel(const PlocState* __restrict__ st, const float4* __restrict__ lo,
                                                           const float4* __restrict__ hi, int* __restrict__ nn) {
    __shared__ float4 slo[PLOC_TPB + 2 * PLOC_RADIUS], shi[PLOC_TPB + 2 * PLOC_RADIUS];
    const int m = st->m;
    for (int base = blockIdx.x * PLOC_TPB; base < m; base += gridDim.x * PLOC_TPB) {
        __syncthreads();
        for (int t = threadIdx.x; t < PLOC_TPB + 2 * PLOC_RADIUS; t += PLOC_TPB) {
            const int g = base - PLOC_RADIUS + t;
            if (g >= 0 && g < m) { slo[t] = lo[g]; shi[t] = hi[g]; }
        }
        __syncthreads();
        const int i = base + threadIdx.x;
        if (i < m) {
            const int c = threadIdx.x + PLOC_RADIUS;
            const float4 alo = slo[c], ahi = shi[c];
            float best = 3.0e38f; int bj = -1;
#pragma unroll
            for (int off = -PLOC_RADIUS; off <= PLOC_RADIUS; off++) {
                if (off == 0) continue;
                const int j = i + off;
                if (j < 0 || j >= m) continue;
                const float a = merged_area(alo, ahi, slo[c + off], shi[c + off]);
                if (a < best) { best = a; bj = j; }   // strict <: ties keep the lowest index
            }
            nn[i] = bj;
        }
    }
}

// flag[i] = (survives << 32) | merges-as-left; zero beyond m (the scan runs over a host-side upper bound of m)
__global__ void __launch_bounds__(PLOC_TPB) ploc_flags_kernel(const PlocState* __restrict__ st, const int* __restrict__ nn,
                                                              unsigned long long* __restrict__ flag, int m_bound) {
    const int m = st->m;
    for (int i = blockIdx.x * PLOC_TPB + threadIdx.x; i < m_bound; i += gridDim.x * PLOC_TPB) {
        unsigned long long f = 0ull;
        if (i < m) {
            const int j = nn[i];
            const bool mutual = j >= 0 && nn[j] == i;
            f = ((unsigned long long)(!(mutual && i > j)) << 32) | (unsigned long long)(mutual && i < j);
        }
        flag[i] = f;
    }
}

__global__ void __launch_bounds__(PLOC_TPB) ploc_scatter_kernel(const PlocState* __restrict__ st, PlocState* __restrict__ st_next,
                                                                const int* __restrict__ nn, const unsigned long long* __restrict__ flag,
                                                                const unsigned long long* __restrict__ scan, const float4* __restrict__ lo,
                                                                const float4* __restrict__ hi, float4* __restrict__ lo_out,
                                                                float4* __restrict__ hi_out, BvhNode* __restrict__ nodes) {
    const int m = st->m, node_base = st->node_base;
    for (int i = blockIdx.x * PLOC_TPB + threadIdx.x; i < m; i += gridDim.x * PLOC_TPB) {
        const unsigned long long f = flag[i], s = scan[i];
        const bool survive = (f >> 32) != 0ull, left = (f & 1ull) != 0ull;
        if (i == m - 1) { st_next->m = (int)(s >> 32) + (int)survive; st_next->node_base = node_base + (int)(s & 0xffffffffull) + (int)left; }
        if (!survive) continue;
        const int pos = (int)(s >> 32);
        float4 alo = lo[i], ahi = hi[i];
        if (left) {
            const int j = nn[i];
            const float4 blo = lo[j], bhi = hi[j];
            const int id = node_base + (int)(s & 0xffffffffull);
            write_node(nodes, id, alo, ahi, blo, bhi);
            alo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), __int_as_float(id));
            ahi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.0f);
        }
        lo_out[pos] = alo; hi_out[pos] = ahi;
    }
}

// ---- single-CTA finish: <= PLOC_FINISH clusters, one per thread, everything in shared memory -----------------------
__global__ void __launch_bounds__(PLOC_FINISH) ploc_finish_kernel(PlocState* st, const float4* __restrict__ lo, const float4* __restrict__ hi,
                                                                  BvhNode* __restrict__ nodes, float* __restrict__ root_box) {
    typedef cub::BlockScan<unsigned long long, PLOC_FINISH> Scan;
    __shared__ typename Scan::TempStorage temp;
    __shared__ float4 slo[PLOC_FINISH], shi[PLOC_FINISH];
    __shared__ int snn[PLOC_FINISH];
    __shared__ int s_m, s_base;
    const int i = threadIdx.x;
    int m = st->m, node_base = st->node_base;
    if (m > PLOC_FINISH) return;   // caller's contract
    if (i < m) { slo[i] = lo[i]; shi[i] = hi[i]; }
    __syncthreads();
    while (m > 1) {
        float4 alo = make_float4(0.f, 0.f, 0.f, 0.f), ahi = alo;
        int bj = -1;
        if (i < m) {
            alo = slo[i]; ahi = shi[i];
            float best = 3.0e38f;
            for (int off = -PLOC_RADIUS; off <= PLOC_RADIUS; off++) {
                const int j = i + off;
                if (off == 0 || j < 0 || j >= m) continue;
                const float a = merged_area(alo, ahi, slo[j], shi[j]);
                if (a < best) { best = a; bj = j; }
            }
            snn[i] = bj;
        }
        __syncthreads();
        bool survive = false, left = false;
        float4 blo = alo, bhi = ahi;
        if (i < m) {
            const bool mutual = bj >= 0 && snn[bj] == i;
            survive = !(mutual && i > bj); left = mutual && i < bj;
            if (left) { blo = slo[bj]; bhi = shi[bj]; }
        }
        const unsigned long long f = ((unsigned long long)survive << 32) | (unsigned long long)left;
        unsigned long long s, total;
        Scan(temp).ExclusiveSum(f, s, total);
        __syncthreads();   // everyone has read slo/shi/snn; temp can be reused next round
        if (survive) {
            const int pos = (int)(s >> 32);
            if (left) {
                const int id = node_base + (int)(s & 0xffffffffull);
                write_node(nodes, id, alo, ahi, blo, bhi);
                alo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), __int_as_float(id));
                ahi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.0f);
            }
            slo[pos] = alo; shi[pos] = ahi;
        }
        if (i == 0) { s_m = (int)(total >> 32); s_base = node_base + (int)(total & 0xffffffffull); }
        __syncthreads();
        m = s_m; node_base = s_base;
    }
    if (i == 0) {
        st->m = m; st->node_base = node_base;
        root_box[0] = slo[0].x; root_box[1] = slo[0].y; root_box[2] = slo[0].z;
        root_box[3] = shi[0].x; root_box[4] = shi[0].y; root_box[5] = shi[0].z;
    }
}

// ---- host ---------------------------------------------------------------------------------------------------
size_t ploc_scan_temp_bytes(int n) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr, n);
    return bytes;
}

// Builds sb.nodes2 (n-1 nodes, root = n-2) from the leaf boxes sb.cl_lo/cl_hi[0] written by K3.  Synchronises the
// stream every few iterations to read the cluster count (the iteration count is data dependent).
int ploc_build(SceneBuffers& sb, cudaStream_t stream, int* launches) {
    const int n = sb.n_faces;
    if (n < 2) return 0;
    cudaError_t e;
#define PL_CU(call) do { e = (call); if (e != cudaSuccess) return (int)e; } while (0)
    const PlocState init = {n, 0};
    PL_CU(cudaMemcpyAsync(sb.ploc_state, &init, sizeof(init), cudaMemcpyHostToDevice, stream));
    int nl = 0, cur = 0, m_bound = n;
    const int grid_cap = 148 * 8;
    PlocState* st = (PlocState*)sb.ploc_state;   // st[0], st[1]: ping-pong with the cluster arrays
    int which = 0;
    while (m_bound > PLOC_FINISH) {
        for (int rep = 0; rep < 4; rep++) {
            const int blocks = min(grid_cap, (m_bound + PLOC_TPB - 1) / PLOC_TPB);
            ploc_nn_kernel<<<blocks, PLOC_TPB, 0, stream>>>(st + which, sb.cl_lo[cur], sb.cl_hi[cur], sb.ploc_nn);
            ploc_flags_kernel<<<blocks, PLOC_TPB, 0, stream>>>(st + which, sb.ploc_nn, sb.ploc_flag, m_bound);
            size_t tb = sb.ploc_temp_bytes;
            PL_CU(cub::DeviceScan::ExclusiveSum(sb.ploc_temp, tb, sb.ploc_flag, sb.ploc_scan, m_bound, stream));
            ploc_scatter_kernel<<<blocks, PLOC_TPB, 0, stream>>>(st + which, st + (which ^ 1), sb.ploc_nn, sb.ploc_flag, sb.ploc_scan,
                                                                 sb.cl_lo[cur], sb.cl_hi[cur], sb.cl_lo[cur ^ 1], sb.cl_hi[cur ^ 1], sb.nodes2);
            nl += 5; cur ^= 1; which ^= 1;
        }
        PlocState h;
        PL_CU(cudaMemcpyAsync(&h, st + which, sizeof(h), cudaMemcpyDeviceToHost, stream));
        PL_CU(cudaStreamSynchronize(stream));
        m_bound = h.m;
    }
    ploc_finish_kernel<<<1, PLOC_FINISH, 0, stream>>>(st + which, sb.cl_lo[cur], sb.cl_hi[cur], sb.nodes2, sb.root_box);
    nl++;
    if (launches) *launches += nl;
    return (int)cudaGetLastError();
#undef PL_CU
}

}  // namespace mcpt
