// ploc.cu — quality pass of seam 1: re-cluster the Morton-ordered leaves into the traversal tree (K4a).
//
// The Karras radix tree (lbvh.cu K2/K3) splits wherever the Morton codes differ; on cornell-box a ray visits 30 of its
// nodes where a surface-area-guided tree needs 12 (measured: DESIGN.md "tree quality"), and traversal — not the build —
// is 99.9 % of a render.  This pass builds the tree that traversal actually walks by parallel locally-ordered
// clustering (Meister & Bittner 2018) over the SAME Morton-sorted leaf sequence the radix tree is built on:
//   repeat until one cluster is left:
//     every cluster looks at its RADIUS neighbours on either side in the (Morton-ordered) cluster array and picks the one
//     whose merged box has the smallest surface area (ties: the buddy i^1, then the lowest index — a mutual pair always exists);
//     mutually-nearest pairs (i < j) become a new internal node at i, j dies;
//     an order-preserving compaction gives every survivor its new position and every merge its node index in array order
//     (so the tree does not depend on scheduling) and writes the new nodes (64-B BvhNode: both child boxes + references).
// The rounds above 1024 clusters run in ONE cooperatively launched persistent kernel (two grid barriers per round), then a
// single CTA finishes the job in shared memory.  Node boxes are exact unions of the
// padded FP32 leaf boxes of K3, so no refit is needed.  Leaf order, leaf indices and the Morton codes are untouched.
#include <cstdio>
#include <cstdlib>
#include <cooperative_groups.h>
#include <cub/block/block_scan.cuh>
#include "lbvh.h"

namespace cg = cooperative_groups;

namespace mcpt {

#ifndef PLOC_RADIUS
#define PLOC_RADIUS 8
#endif
#define PLOC_GRID_TPB 256      /* cooperative-grid variant of the rounds kernel */
#define PLOC_CLUSTER_TPB 1024  /* one-cluster variant */
#define PLOC_CLUSTER_CTAS 8
#define PLOC_CLUSTER_MAX_N 0   /* measured: cornell-box 125 us in one cluster (8 SMs do all the work) against 91 us on the cooperative grid (59 SMs): parallelism beats barrier latency; the variant stays for A/B.  Also measured: the grid down to 4096 clusters, then one cluster of 8 x 256 threads down to 1024: 38 + 62 us against 84 us — two tiles per CTA and round with 8 warps per SM hide no latency */
#define PLOC_FINISH 1024
#define PLOC_MAX_ITERS 4096

__device__ __forceinline__ float merged_area(const float4 alo, const float4 ahi, const float4 blo, const float4 bhi) {
    const float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y),
                dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dy), __fmul_rn(dy, dz)), __fmul_rn(dz, dx));   // symmetric in (a, b), no FMA
}

// Nearest neighbour of cluster i among i-RADIUS .. i+RADIUS (global indices, 0 <= j < m): smallest merged area; among equal
// areas the buddy i^1 first, then the lowest index.  A mutual pair always exists: take the globally smallest area A*; if some
// buddy pair (i, i^1) reaches A*, both pick each other; otherwise the preference never applies at A* and the lowest-index
// cluster taking part in an A* pair and its lowest-index A* partner pick each other.  The buddy rule is what keeps
// tie-heavy input (regular grids, duplicated triangles: every area equal) at n/2 merges per round instead of ONE.
// slo/shi hold the boxes of clusters base .. ; `at` = index of cluster i in them.
__device__ __forceinline__ int ploc_nearest(const float4* __restrict__ slo, const float4* __restrict__ shi, const int at, const int i, const int m) {
    const float4 alo = slo[at], ahi = shi[at];
    float best = 3.0e38f; int bj = -1;
#pragma unroll
    for (int off = -PLOC_RADIUS; off <= PLOC_RADIUS; off++) {
        if (off == 0) continue;
        const int j = i + off;
        // no branch on the range test: the sixteen box loads are independent and issue back to back (slo/shi extend RADIUS entries
        // beyond the clusters on either side, so an out-of-range neighbour reads stale but addressable memory and is masked)
        const bool in_range = j >= 0 && j < m;
        const float a = merged_area(alo, ahi, slo[at + off], shi[at + off]);
        if (in_range && (a < best || (a == best && j == (i ^ 1)))) { best = a; bj = j; }   // strict <: ties keep the lowest index, but for the buddy
    }
    return bj;
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

// ---- the rounds above PLOC_FINISH clusters: ONE persistent, cooperatively launched kernel -----------------------------
// (Round 1 ran every round as four kernels + a CUB scan and synchronised with the host every fourth round to read the
// cluster count: ~60 launches and three host round trips for cornell-box, 0.36 of the build's 0.60 ms.)  Every CTA owns a
// contiguous range of the cluster array, so the order-preserving compaction needs only per-CTA counts:
//   phase 1  per cluster of a 256-cluster tile, and per cluster of its RADIUS-wide rims: nearest neighbour, into shared memory
//            (boxes staged with a 2 x RADIUS halo) — the mutual test reads the partner's choice there, no second pass over
//            global memory; one flag byte per cluster (survives | merges-as-left | offset of the partner), per-CTA counts of
//            survivors and merges                                                                      -> grid.sync
//   phase 2  every CTA sums the counts of the CTAs before it (and all of them: the next round's cluster count, identical
//            in every CTA, no global state), scans its flags tile by tile and scatters survivors / writes the new nodes
//            in array order — so node numbering does not depend on scheduling                          -> grid.sync
#define PLOC_HALO (2 * PLOC_RADIUS)
struct PlocState { int m; int node_base; int error; int rounds; int finish_rounds; };   // clusters alive, internal nodes created, stalled, rounds run (grid kernel | buffer parity << 30), rounds of the finish kernel

// CLUSTER: the CTAs are ONE thread-block cluster (8 CTAs of 1024 threads) and the barrier is the hardware cluster barrier — a
// fraction of a microsecond against ~3 us for a grid-wide barrier through L2; used up to 64 Ki triangles, where a round is
// nothing but barrier latency (cornell-box: 7 rounds).  Otherwise: a cooperative grid that fills the GPU.
template <int PLOC_TPB, bool CLUSTER>
__global__ void __launch_bounds__(PLOC_TPB) ploc_rounds_kernel(PlocState* st, float4* lo0, float4* hi0, float4* lo1, float4* hi1,
                                                               unsigned char* __restrict__ flagb, uint2* blockcnt, BvhNode* __restrict__ nodes,
                                                               const int n) {
    auto all_ctas_sync = [] {
        if (CLUSTER) cg::this_cluster().sync(); else cg::this_grid().sync();
    };
    typedef cub::BlockScan<unsigned long long, PLOC_TPB> Scan;
    __shared__ typename Scan::TempStorage temp;
    __shared__ float4 slo[PLOC_TPB + 2 * PLOC_HALO], shi[PLOC_TPB + 2 * PLOC_HALO];
    __shared__ int snn[PLOC_TPB + 2 * PLOC_RADIUS];
    __shared__ unsigned s_red[2][PLOC_TPB / 32];
    __shared__ unsigned long long s_pre[2];
    const int G = gridDim.x, b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int m = n, node_base = 0, rounds = 0, error = 0;
    float4 *lo = lo0, *hi = hi0, *lo_out = lo1, *hi_out = hi1;
    while (m > PLOC_FINISH) {
        const int per = ((m + G - 1) / G + PLOC_TPB - 1) / PLOC_TPB * PLOC_TPB;
        const int r0 = min(b * per, m), r1 = min(r0 + per, m);
        // ---- phase 1 ----
        unsigned n_surv = 0, n_left = 0;
        for (int base = r0; base < r1; base += PLOC_TPB) {
            __syncthreads();
            for (int t = tid; t < PLOC_TPB + 2 * PLOC_HALO; t += PLOC_TPB) {
                const int g = base - PLOC_HALO + t;
                if (g >= 0 && g < m) { slo[t] = lo[g]; shi[t] = hi[g]; }
            }
            __syncthreads();
            // nearest neighbour of every cluster of the tile AND of the RADIUS clusters on either side (whose own neighbourhoods
            // still lie inside the 2 x RADIUS halo): the mutual test then reads its partner's choice from shared memory
            const int i = base + tid;
            if (i < m) snn[tid + PLOC_RADIUS] = ploc_nearest(slo, shi, tid + PLOC_HALO, i, m);
            if (tid < 2 * PLOC_RADIUS) {
                const int t = tid < PLOC_RADIUS ? tid : PLOC_TPB + tid;          // snn index: [0, R) left halo, [TPB+R, TPB+2R) right halo
                const int g = base - PLOC_RADIUS + t;
                if (g >= 0 && g < m) snn[t] = ploc_nearest(slo, shi, t + PLOC_RADIUS, g, m);
            }
            __syncthreads();
            bool survive = false, left = false;
            if (i < r1) {
                const int j = snn[tid + PLOC_RADIUS];
                const bool mutual = j >= 0 && snn[j - base + PLOC_RADIUS] == i;
                survive = !(mutual && i > j); left = mutual && i < j;
                flagb[i] = (unsigned char)((unsigned)survive | ((unsigned)left << 1) | ((unsigned)(j - i + PLOC_RADIUS) << 2));
            }
            const unsigned ms = __ballot_sync(0xffffffffu, survive), ml = __ballot_sync(0xffffffffu, left);
            if (lane == 0) { n_surv += __popc(ms); n_left += __popc(ml); }
        }
        if (lane == 0) { s_red[0][warp] = n_surv; s_red[1][warp] = n_left; }
        __syncthreads();
        if (tid == 0) {
            unsigned a = 0, c = 0;
            for (int w = 0; w < PLOC_TPB / 32; w++) { a += s_red[0][w]; c += s_red[1][w]; }
            blockcnt[b] = make_uint2(a, c);
        }
        all_ctas_sync();
        // ---- phase 2 ----
        unsigned long long before = 0ull, total = 0ull;   // packed (survivors << 32 | merges)
        for (int k = tid; k < G; k += PLOC_TPB) {
            const uint2 c = __ldcg(&blockcnt[k]);
            const unsigned long long v = ((unsigned long long)c.x << 32) | c.y;
            total += v; if (k < b) before += v;
        }
        before = __reduce_add_sync(0xffffffffu, (unsigned)(before >> 32)) * (1ull << 32) + __reduce_add_sync(0xffffffffu, (unsigned)before);
        total = __reduce_add_sync(0xffffffffu, (unsigned)(total >> 32)) * (1ull << 32) + __reduce_add_sync(0xffffffffu, (unsigned)total);
        __syncthreads();   // s_red free again
        if (lane == 0) { s_red[0][warp] = (unsigned)(before >> 32); s_red[1][warp] = (unsigned)before; }
        __syncthreads();
        if (tid == 0) {
            unsigned a = 0, c = 0;
            for (int w = 0; w < PLOC_TPB / 32; w++) { a += s_red[0][w]; c += s_red[1][w]; }
            s_pre[0] = ((unsigned long long)a << 32) | c;
        }
        __syncthreads();
        if (lane == 0) { s_red[0][warp] = (unsigned)(total >> 32); s_red[1][warp] = (unsigned)total; }
        unsigned long long running = s_pre[0];
        __syncthreads();
        if (tid == 0) {
            unsigned a = 0, c = 0;
            for (int w = 0; w < PLOC_TPB / 32; w++) { a += s_red[0][w]; c += s_red[1][w]; }
            s_pre[1] = ((unsigned long long)a << 32) | c;
        }
        __syncthreads();
        total = s_pre[1];
        for (int base = r0; base < r1; base += PLOC_TPB) {
            const int i = base + tid;
            unsigned f = 0u;
            if (i < r1) f = flagb[i];
            const bool survive = f & 1u, left = (f >> 1) & 1u;
            unsigned long long s, tile_total;
            Scan(temp).ExclusiveSum(((unsigned long long)survive << 32) | (unsigned long long)left, s, tile_total);
            __syncthreads();   // temp reusable next tile
            s += running; running += tile_total;
            if (survive) {
                const int pos = (int)(s >> 32);
                float4 alo = lo[i], ahi = hi[i];
                if (left) {
                    const int j = i + (int)(f >> 2) - PLOC_RADIUS;
                    const float4 blo = lo[j], bhi = hi[j];
                    const int id = node_base + (int)(s & 0xffffffffull);
                    write_node(nodes, id, alo, ahi, blo, bhi);
                    alo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), __int_as_float(id));
                    ahi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.0f);
                }
                lo_out[pos] = alo; hi_out[pos] = ahi;
            }
        }
        const int m_next = (int)(total >> 32);
        node_base += (int)(total & 0xffffffffull);
        rounds++;
        if (m_next >= m || rounds >= PLOC_MAX_ITERS) { error = 1; break; }   // no mutual pair (cannot happen on finite boxes) or runaway
        m = m_next;
        float4* t = lo; lo = lo_out; lo_out = t; t = hi; hi = hi_out; hi_out = t;
        all_ctas_sync();
    }
    if (b == 0 && tid == 0) { st->m = m; st->node_base = node_base; st->error = error; st->rounds = rounds | ((lo == lo0 ? 0 : 1) << 30); }
}

// ---- single-CTA finish: <= PLOC_FINISH clusters, one per thread, everything in shared memory -----------------------
__global__ void __launch_bounds__(PLOC_FINISH) ploc_finish_kernel(PlocState* st, const float4* __restrict__ lo0, const float4* __restrict__ hi0,
                                                                  const float4* __restrict__ lo1, const float4* __restrict__ hi1,
                                                                  BvhNode* __restrict__ nodes, float* __restrict__ root_box, unsigned* w4_ctr) {
    __shared__ float4 slo[PLOC_FINISH + 2 * PLOC_RADIUS], shi[PLOC_FINISH + 2 * PLOC_RADIUS];   // clusters at [RADIUS, RADIUS + m)
    __shared__ int snn[PLOC_FINISH];
    __shared__ unsigned s_cnt[1][32];
    const int i = threadIdx.x, lane = i & 31, warp = i >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    if (i == 0 && w4_ctr) { w4_ctr[0] = 1u; w4_ctr[1] = 0u; w4_ctr[2] = 0u; w4_ctr[3] = 0u; w4_ctr[4] = 0u; }   // wide.cu's W4_* counters
    int m = st->m, node_base = st->node_base;
    if (m > PLOC_FINISH || st->error) return;   // the rounds kernel stalled: mcpt_build_bvh falls back to the radix tree
    const bool second = (st->rounds >> 30) & 1;
    const float4* lo = second ? lo1 : lo0; const float4* hi = second ? hi1 : hi0;
    if (i < m) { slo[i + PLOC_RADIUS] = lo[i]; shi[i + PLOC_RADIUS] = hi[i]; }
    __syncthreads();
    int rounds = 0;
    // Only the warps that still hold clusters take part in a round: a warp whose first thread is beyond the cluster count leaves
    // for good (the count only shrinks) and the others meet at a NAMED barrier sized for them — with all 32 warps at every barrier
    // and every warp repeating the bookkeeping, two thirds of the kernel's 177 k warp instructions were issued by warps with
    // nothing to cluster, and `barrier` was its top stall (13 warps per issue).
    while (m > 1) {
        const int n_act = (m + 31) >> 5;                      // warps with clusters this round; the same value in every thread
        if (warp >= n_act) return;
        const int bar_threads = n_act * 32;
        auto act_sync = [&] { asm volatile("bar.sync 1, %0;" :: "r"(bar_threads) : "memory"); };
        float4 alo = make_float4(0.f, 0.f, 0.f, 0.f), ahi = alo;
        int bj = -1;
        if (i < m) {
            alo = slo[i + PLOC_RADIUS]; ahi = shi[i + PLOC_RADIUS];
            bj = ploc_nearest(slo, shi, i + PLOC_RADIUS, i, m);
            snn[i] = bj;
        }
        act_sync();
        bool survive = false, left = false;
        float4 blo = alo, bhi = ahi;
        if (i < m) {
            const bool mutual = bj >= 0 && snn[bj] == i;
            survive = !(mutual && i > bj); left = mutual && i < bj;
            if (left) { blo = slo[bj + PLOC_RADIUS]; bhi = shi[bj + PLOC_RADIUS]; }
        }
        // order-preserving compaction with warp ballots; the per-warp counts (survivors | merges << 16, both <= 1024) are summed by
        // every active warp with two REDUX instructions — all counts, and the counts of the warps before it — so the totals need no
        // broadcast: three barriers per round where a 1024-thread cub::BlockScan took six
        const unsigned ms = __ballot_sync(0xffffffffu, survive), ml = __ballot_sync(0xffffffffu, left);
        if (lane == 0) s_cnt[0][warp] = (unsigned)__popc(ms) | ((unsigned)__popc(ml) << 16);
        act_sync();   // everyone has read slo/shi/snn; the counts are in
        const unsigned c = lane < n_act ? s_cnt[0][lane] : 0u;
        const unsigned tot = __reduce_add_sync(0xffffffffu, c), before = __reduce_add_sync(0xffffffffu, lane < warp ? c : 0u);
        const int total_s = (int)(tot & 0xffffu), total_l = (int)(tot >> 16);
        const int off_s = (int)(before & 0xffffu), off_l = (int)(before >> 16);
        if (survive) {
            const int pos = off_s + __popc(ms & lt_mask);
            if (left) {
                const int id = node_base + off_l + __popc(ml & lt_mask);
                write_node(nodes, id, alo, ahi, blo, bhi);
                alo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), __int_as_float(id));
                ahi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.0f);
            }
            slo[pos + PLOC_RADIUS] = alo; shi[pos + PLOC_RADIUS] = ahi;
        }
        act_sync();
        if (total_s >= m || ++rounds > PLOC_MAX_ITERS) { if (i == 0) st->error = 1; return; }
        m = total_s; node_base += total_l;
    }
    if (i == 0) {
        st->m = m; st->node_base = node_base; st->finish_rounds = rounds;
        root_box[0] = slo[PLOC_RADIUS].x; root_box[1] = slo[PLOC_RADIUS].y; root_box[2] = slo[PLOC_RADIUS].z;
        root_box[3] = shi[PLOC_RADIUS].x; root_box[4] = shi[PLOC_RADIUS].y; root_box[5] = shi[PLOC_RADIUS].z;
    }
}

// ---- host ---------------------------------------------------------------------------------------------------
size_t ploc_scratch_bytes(int n) {   // per-CTA counts of the rounds kernel (<= 8 CTAs per SM on any device we accept)
    (void)n;
    return sizeof(uint2) * 4096;
}

// Builds sb.nodes2 (n-1 nodes, root = n-2) from the leaf boxes sb.cl_lo/cl_hi[0] written by K3: two launches, no host
// round trip.  A stall (st->error, checked by mcpt_build_bvh after the build) makes the caller fall back to the radix tree.
// init_w4: let the finish kernel also reset wide.cu's counters (saves wide_build its own init launch).
int ploc_build(SceneBuffers& sb, cudaStream_t stream, int* launches, bool init_w4) {
    const int n = sb.n_faces;
    if (n < 2) return 0;
    cudaError_t e;
    PlocState* st = (PlocState*)sb.ploc_state;
    float4 *lo0 = sb.cl_lo[0], *hi0 = sb.cl_hi[0], *lo1 = sb.cl_lo[1], *hi1 = sb.cl_hi[1];
    unsigned char* flagb = sb.ploc_flag; uint2* blockcnt = (uint2*)sb.ploc_temp; BvhNode* nodes = sb.nodes2; int nn = n;
    if (n <= PLOC_CLUSTER_MAX_N) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(PLOC_CLUSTER_CTAS); cfg.blockDim = dim3(PLOC_CLUSTER_TPB); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = PLOC_CLUSTER_CTAS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        e = cudaLaunchKernelEx(&cfg, ploc_rounds_kernel<PLOC_CLUSTER_TPB, true>, st, lo0, hi0, lo1, hi1, flagb, blockcnt, nodes, nn);
        if (e != cudaSuccess) return (int)e;
    } else {
        if (!sb.ploc_resident_blocks) {
            int dev = 0, sms = 0, per_sm = 0, coop = 0;
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ploc_rounds_kernel<PLOC_GRID_TPB, false>, PLOC_GRID_TPB, 0);
            if (e != cudaSuccess || sms <= 0 || per_sm <= 0 || !coop) return (int)(e != cudaSuccess ? e : cudaErrorInvalidConfiguration);
            sb.ploc_resident_blocks = min(sms * per_sm, 4096);
        }
        int blocks = min(sb.ploc_resident_blocks, (n + PLOC_GRID_TPB - 1) / PLOC_GRID_TPB);
        void* args[] = {(void*)&st, (void*)&lo0, (void*)&hi0, (void*)&lo1, (void*)&hi1, (void*)&flagb, (void*)&blockcnt, (void*)&nodes, (void*)&nn};
        e = cudaLaunchCooperativeKernel((const void*)ploc_rounds_kernel<PLOC_GRID_TPB, false>, dim3(blocks), dim3(PLOC_GRID_TPB), args, 0, stream);
        if (e != cudaSuccess) return (int)e;
    }
    ploc_finish_kernel<<<1, PLOC_FINISH, 0, stream>>>(st, lo0, hi0, lo1, hi1, sb.nodes2, sb.root_box, init_w4 ? sb.w4_ctr : nullptr);
    if (launches) *launches += 2;
    return (int)cudaGetLastError();
}

bool ploc_ok(const SceneBuffers& sb, cudaError_t* err) {
    PlocState h{};
    *err = cudaMemcpy(&h, sb.ploc_state, sizeof(h), cudaMemcpyDeviceToHost);
    if (getenv("MCPT_DIAG_PLOC")) std::fprintf(stderr, "[diag] PLOC: %d grid rounds + %d finish rounds\n", h.rounds & 0x3fffffff, h.finish_rounds);
    return *err == cudaSuccess && h.error == 0 && h.m == 1;
}

}  // namespace mcpt
