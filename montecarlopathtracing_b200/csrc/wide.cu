// wide.cu — K4b: collapse the binary traversal tree into a 4-wide tree of 64-B quantised nodes (QNode4).
//
// Why: wf_trace is bound by the latency of one dependent node fetch per step with ~36 resident warps per SM (ncu: issue
// slots half empty, long-scoreboard the top stall).  A 4-wide node halves the number of dependent steps per ray and gives
// every step four independent slab tests, and the node array shrinks from (n-1) x 32 B to ~(n/2.4) x 64 B.
//
// How: breadth-first, ONE launch of a fully resident grid.  A thread owns the wide nodes t, t + stride, ... and takes them in
// order: it waits until queue[t] names the binary node the wide node is rooted at, opens that node's two children, then twice more the open child with the largest surface area (up to four
// children; a bottom node over two single triangles counts as a leaf, see fold_leaf_pair), reserves consecutive wide-node
// indices for its internal children with one atomicAdd on the queue tail and writes their binary roots into the queue.
// Producers always have a lower index than their consumers, every thread is resident and walks its entries in increasing
// order, so the lowest unprocessed entry is always one some thread is polling right now: the wait cannot deadlock.  A warp
// polls with all its lanes in step (no lane spins while a lane of the same warp that it depends on is held back), backs off
// with __nanosleep when nothing arrived, and the wait is bounded (error flag instead of a hang).
// The layout is breadth-first, so the top of the tree is contiguous; which sibling subtree gets the lower index depends on
// the order of the atomics, the TREE (and therefore every traversal result) does not.
#include <cooperative_groups.h>
#include "lbvh.h"
#include "quant.cuh"

namespace cg = cooperative_groups;

namespace mcpt {

// W4_LEVELS = number of internal levels of the wide tree (root = level 1): what bounds the traversal stack (see wide_build)
enum { W4_TAIL = 0, W4_PROCESSED = 1, W4_FINISHED = 2, W4_ERROR = 3, W4_LEVELS = 4 };
#define W4_EMPTY (-1)
// a queue entry = binary node id | (level of the wide node - 1) << 26; levels saturate at 31 (any tree that deep is rejected)
#define W4_ID_MASK ((1 << MCPT_LEAF_BITS) - 1)
#define W4_SPIN_LIMIT (1 << 20)

struct W4Slot { float lo[3], hi[3]; int ref; };

__device__ __forceinline__ void w4_children(const BvhNode* __restrict__ nodes, const int id, const bool fold, W4Slot& a, W4Slot& b) {
    const float4 q0 = nodes[id].q[0], q1 = nodes[id].q[1], q2 = nodes[id].q[2], q3 = nodes[id].q[3];
    a.lo[0] = q0.x; a.hi[0] = q0.y; a.lo[1] = q0.z; a.hi[1] = q0.w; a.lo[2] = q2.x; a.hi[2] = q2.y;
    b.lo[0] = q1.x; b.hi[0] = q1.y; b.lo[1] = q1.z; b.hi[1] = q1.w; b.lo[2] = q2.z; b.hi[2] = q2.w;
    a.ref = fold_leaf_pair(nodes, __float_as_int(q3.x), fold);
    b.ref = fold_leaf_pair(nodes, __float_as_int(q3.y), fold);
}
__device__ __forceinline__ float w4_area(const W4Slot& s) {
    const float dx = s.hi[0] - s.lo[0], dy = s.hi[1] - s.lo[1], dz = s.hi[2] - s.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

__global__ void __launch_bounds__(256) wide_init_kernel(unsigned* ctr) {
    ctr[W4_TAIL] = 1u; ctr[W4_PROCESSED] = 0u; ctr[W4_FINISHED] = 0u; ctr[W4_ERROR] = 0u; ctr[W4_LEVELS] = 0u;
}

__global__ void __launch_bounds__(256) wide_collapse_kernel(const BvhNode* __restrict__ nodes, const int root2, const int max_nodes,
                                                            const float* __restrict__ root_box, QNode4* __restrict__ out,
                                                            int* queue, unsigned* ctr, const bool fold, QGrid* __restrict__ qgrid) {
    const int stride = gridDim.x * blockDim.x, lane = threadIdx.x & 31;
    float o[3], cell[3], inv[3];
    quant_grid(root_box, o, cell, inv);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        qgrid->origin = make_float4(o[0], o[1], o[2], 0.0f);
        qgrid->cell = make_float4(cell[0], cell[1], cell[2], 0.0f);
    }
    unsigned deepest = 0u;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x;; t += stride) {   // warp-uniform trip count: the lanes of a warp advance together
        bool done = t >= max_nodes;
        bool fin = false;
        for (int spin = 0;; spin++) {
            int id = W4_EMPTY;
            if (!done) id = t == 0 ? root2 : *(volatile int*)&queue[t];
            const bool ready = !done && id != W4_EMPTY;
            const unsigned rm = __ballot_sync(0xffffffffu, ready);
            if (rm != 0u) {
                // the lanes whose entry has arrived collapse their node; ONE tail reservation and ONE processed update per warp
                // (same-address atomics from every thread were the bound of the first version: 2.5 ms for 5 M triangles)
                W4Slot s[4];
                int cnt = 0, m = 0;
                int level = 0;
                if (ready) {
                    level = t == 0 ? 0 : (id >> MCPT_LEAF_BITS);
                    id &= W4_ID_MASK;
                    deepest = max(deepest, (unsigned)level + 1u);
                    cnt = 2;
                    w4_children(nodes, id, fold, s[0], s[1]);
                    for (int step = 0; step < 2; step++) {
                        int pick = -1; float best = -1.0f;
                        for (int j = 0; j < cnt; j++) {
                            const float a = w4_area(s[j]);
                            if (s[j].ref >= 0 && a > best) { best = a; pick = j; }
                        }
                        if (pick < 0) break;
                        const int open = s[pick].ref;
                        w4_children(nodes, open, fold, s[pick], s[cnt]);
                        cnt++;
                    }
                    for (int j = 0; j < cnt; j++) m += s[j].ref >= 0;
                }
                int incl = m;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
                const int total = __shfl_sync(0xffffffffu, incl, 31);
                unsigned base = 0;
                if (lane == 0 && total) base = atomicAdd(&ctr[W4_TAIL], (unsigned)total);
                base = __shfl_sync(0xffffffffu, base, 0) + (unsigned)(incl - m);
                if (ready) {
                    QNode4 q;
                    int k = 0;
                    for (int j = 0; j < 4; j++) {
                        if (j < cnt) {
                            int ref = s[j].ref;
                            if (ref >= 0) { const int w = (int)base + k++; *(volatile int*)&queue[w] = ref | (min(level + 1, 31) << MCPT_LEAF_BITS); ref = w; }
                            q.c[j] = make_uint4(quant_lo(s[j].lo[0], o[0], inv[0]) | (quant_hi(s[j].hi[0], o[0], inv[0]) << 16),
                                                quant_lo(s[j].lo[1], o[1], inv[1]) | (quant_hi(s[j].hi[1], o[1], inv[1]) << 16),
                                                quant_lo(s[j].lo[2], o[2], inv[2]) | (quant_hi(s[j].hi[2], o[2], inv[2]) << 16), (unsigned)ref);
                        } else {
                            q.c[j] = make_uint4(0x0000FFFFu, 0x0000FFFFu, 0x0000FFFFu, 0x80000000u);   // inverted box: never hit
                        }
                    }
                    out[t] = q;
                    done = true;
                }
                __threadfence();
                __syncwarp();
                if (lane == 0) {
                    const unsigned p = atomicAdd(&ctr[W4_PROCESSED], (unsigned)__popc(rm)) + (unsigned)__popc(rm);
                    if (p == atomicAdd(&ctr[W4_TAIL], 0u)) *(volatile unsigned*)&ctr[W4_FINISHED] = 1u;
                }
            }
            const bool all_done = __all_sync(0xffffffffu, done);
            if (all_done) break;
            if ((spin & 3) == 3) {   // nothing new for this warp: look at the finished flag, then back off a little
                fin = *(volatile unsigned*)&ctr[W4_FINISHED] != 0u;
                if (fin) break;
                __nanosleep(64);
            }
            if (spin > W4_SPIN_LIMIT) { ctr[W4_ERROR] = 1u; *(volatile unsigned*)&ctr[W4_FINISHED] = 1u; fin = true; break; }   // releases every waiter
        }
        if (__any_sync(0xffffffffu, fin)) break;
        if (t - lane + stride >= max_nodes) break;   // no entries left for this warp
    }
    deepest = __reduce_max_sync(0xffffffffu, deepest);
    if (lane == 0 && deepest) atomicMax(&ctr[W4_LEVELS], deepest);
}


// ---- small trees: candidates for every binary node at once, then ONE thread-block cluster, level by level ---------------------
// Below ~64 Ki triangles the kernel above is nothing but a latency chain: a level's nodes are found by POLLING the queue through
// L2, one level after the other (cornell-box: 118 us for 6 300 wide nodes in ~10 levels).  Two steps instead:
//   wide_candidates_kernel   what EVERY binary node would look like as the root of a wide node — its (up to) four children by the
//                            opening rule above, boxes already on the 16-bit grid, references still binary ids.  No node depends
//                            on another, so the ~8 dependent L2 round trips of the rule (node, children, folded grandchildren,
//                            twice more) are paid ONCE, by all nodes in parallel, instead of once per level.
//   wide_collapse_small_kernel  the 8 CTAs of one cluster take a whole level at a time — its entries are the queue slots
//                            [begin, end) — read each entry's candidate (one 64-B fetch), number its internal children with one
//                            atomic per warp and meet at the hardware cluster barrier, ONE per level: level L counts its new
//                            entries in counter L mod 3, which is read after the barrier and cleared again during level L + 1.
// Same collapse rule, same tree as the polling kernel; sibling numbering again depends on the order of the atomics only.
// cornell-box: 94 us (one kernel, two barriers and ~8 round trips per level) -> 8 + 35 us; veach-mis' build 0.23 -> 0.17 ms.
#define W4_SMALL_CTAS 8
#define W4_SMALL_TPB 512
enum { W4_LEVEL_CTR = 8 };   // ctr[8..10]: the rotating per-level counters (zero outside the kernel)

__global__ void __launch_bounds__(256) wide_candidates_kernel(const BvhNode* __restrict__ nodes, const int n_nodes, const float* __restrict__ root_box,
                                                              QNode4* __restrict__ cand, const bool fold, QGrid* __restrict__ qgrid) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    float o[3], cell[3], inv[3];
    quant_grid(root_box, o, cell, inv);
    if (id == 0) {
        qgrid->origin = make_float4(o[0], o[1], o[2], 0.0f);
        qgrid->cell = make_float4(cell[0], cell[1], cell[2], 0.0f);
    }
    if (id >= n_nodes) return;
    W4Slot s[4];
    int cnt = 2;
    w4_children(nodes, id, fold, s[0], s[1]);
    for (int step = 0; step < 2; step++) {
        int pick = -1; float best = -1.0f;
        for (int j = 0; j < cnt; j++) {
            const float a = w4_area(s[j]);
            if (s[j].ref >= 0 && a > best) { best = a; pick = j; }
        }
        if (pick < 0) break;
        const int open = s[pick].ref;
        w4_children(nodes, open, fold, s[pick], s[cnt]);
        cnt++;
    }
    QNode4 q;
    for (int j = 0; j < 4; j++) {
        if (j < cnt) {
            q.c[j] = make_uint4(quant_lo(s[j].lo[0], o[0], inv[0]) | (quant_hi(s[j].hi[0], o[0], inv[0]) << 16),
                                quant_lo(s[j].lo[1], o[1], inv[1]) | (quant_hi(s[j].hi[1], o[1], inv[1]) << 16),
                                quant_lo(s[j].lo[2], o[2], inv[2]) | (quant_hi(s[j].hi[2], o[2], inv[2]) << 16), (unsigned)s[j].ref);
        } else {
            q.c[j] = make_uint4(0x0000FFFFu, 0x0000FFFFu, 0x0000FFFFu, 0x80000000u);   // inverted box: never hit
        }
    }
    cand[id] = q;
}

__global__ void __launch_bounds__(W4_SMALL_TPB) wide_collapse_small_kernel(const QNode4* __restrict__ cand, const int root2, const int max_nodes,
                                                                           QNode4* __restrict__ out, int* queue, unsigned* ctr) {
    cg::cluster_group cluster = cg::this_cluster();
    const int nthreads = W4_SMALL_CTAS * W4_SMALL_TPB, gtid = blockIdx.x * W4_SMALL_TPB + threadIdx.x, lane = threadIdx.x & 31;
    int begin = 0, end = 1;
    unsigned levels = 0;
    __shared__ unsigned s_new;
    while (begin < end && end <= max_nodes) {
        unsigned* level_ctr = ctr + W4_LEVEL_CTR + levels % 3u;
        for (int base = begin + (gtid - lane); base < end; base += nthreads) {   // warp-uniform trip count
            const int t = base + lane;
            const bool active = t < end;
            uint4 c0 = make_uint4(0u, 0u, 0u, 0x80000000u), c1 = c0, c2 = c0, c3 = c0;
            int m = 0;
            if (active) {
                const int id = t == 0 ? root2 : __ldcg(&queue[t]);
                c0 = cand[id].c[0]; c1 = cand[id].c[1]; c2 = cand[id].c[2]; c3 = cand[id].c[3];
                m = ((int)c0.w >= 0) + ((int)c1.w >= 0) + ((int)c2.w >= 0) + ((int)c3.w >= 0);
            }
            int incl = m;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
            const int total = __shfl_sync(0xffffffffu, incl, 31);
            unsigned at = 0;
            if (lane == 0 && total) at = atomicAdd(level_ctr, (unsigned)total);
            at = (unsigned)end + __shfl_sync(0xffffffffu, at, 0) + (unsigned)(incl - m);
            if (active) {
                if (at + (unsigned)m <= (unsigned)max_nodes) {   // (a tree can never need more: the test only guards the queue)
                    if ((int)c0.w >= 0) { __stcg(&queue[at], (int)c0.w); c0.w = at++; }
                    if ((int)c1.w >= 0) { __stcg(&queue[at], (int)c1.w); c1.w = at++; }
                    if ((int)c2.w >= 0) { __stcg(&queue[at], (int)c2.w); c2.w = at++; }
                    if ((int)c3.w >= 0) { __stcg(&queue[at], (int)c3.w); c3.w = at++; }
                }
                out[t].c[0] = c0; out[t].c[1] = c1; out[t].c[2] = c2; out[t].c[3] = c3;
            }
        }
        cluster.sync();                                   // the level's queue entries and its count are complete
        if (threadIdx.x == 0) s_new = __ldcg(level_ctr);
        if (gtid == 0) ctr[W4_LEVEL_CTR + (levels + 2u) % 3u] = 0u;   // last read before this barrier, next used after the next one
        __syncthreads();
        begin = end; end += (int)s_new;                   // (s_new is rewritten only after the next cluster barrier)
        levels++;
    }
    cluster.sync();
    if (gtid == 0) {
        ctr[W4_LEVEL_CTR] = 0u; ctr[W4_LEVEL_CTR + 1] = 0u; ctr[W4_LEVEL_CTR + 2] = 0u;
        ctr[W4_TAIL] = (unsigned)end; ctr[W4_PROCESSED] = (unsigned)end; ctr[W4_LEVELS] = levels; ctr[W4_ERROR] = end > max_nodes ? 1u : 0u; ctr[W4_FINISHED] = 1u;
    }
}

int wide_build(SceneBuffers& sb, const BvhNode* walk, int root2, cudaStream_t stream, int* launches, bool prepared) {
    const int n = sb.n_faces;
    const int max_nodes = n > 1 ? n - 1 : 1;
    cudaError_t e = cudaSuccess;
    if (!prepared) {
        e = cudaMemsetAsync(sb.w4_queue, 0xFF, sizeof(int) * (size_t)max_nodes, stream);
        if (e != cudaSuccess) return (int)e;
        wide_init_kernel<<<1, 1, 0, stream>>>(sb.w4_ctr);
        if (launches) *launches += 2;
    }
    if (max_nodes <= W4_SMALL_MAX_N) {   // small trees: candidates for every binary node, then one cluster, level by level
        const bool fold_s = sb.fold_pairs && n > 2;
        wide_candidates_kernel<<<(max_nodes + 255) / 256, 256, 0, stream>>>(walk, max_nodes, (const float*)sb.root_box, sb.w4_cand, fold_s, sb.qgrid);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(W4_SMALL_CTAS); cfg.blockDim = dim3(W4_SMALL_TPB); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = W4_SMALL_CTAS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        e = cudaLaunchKernelEx(&cfg, wide_collapse_small_kernel, (const QNode4*)sb.w4_cand, root2, max_nodes, sb.qnodes4, sb.w4_queue, sb.w4_ctr);
        if (launches) *launches += 2;
        return (int)(e != cudaSuccess ? e : cudaGetLastError());
    }
    // large trees: a fully RESIDENT grid polling the queue (the no-deadlock argument needs every thread to be running): launched
    // COOPERATIVELY, so the runtime refuses the launch instead of letting CTAs wait for SMs that another context or stream holds;
    // the grid size comes from this device's occupancy (queried once per context: sb.w4_resident_blocks).  The opening rule is
    // evaluated on demand here: candidates for all n - 1 binary nodes (3.6 x the nodes that end up in the tree) cost more than the
    // shorter polling chain saves (measured: 1 M triangles 0.98 -> 0.94 ms, 5 M 2.73 -> 2.77 ms; the large case decides).
    if (!sb.w4_resident_blocks) {
        int dev = 0, sms = 0, per_sm = 0, coop = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wide_collapse_kernel, 256, 0);
        if (e != cudaSuccess || sms <= 0 || per_sm <= 0 || !coop) return (int)(e != cudaSuccess ? e : cudaErrorInvalidConfiguration);
        sb.w4_resident_blocks = sms * per_sm;
    }
    int blocks = (max_nodes + 255) / 256;
    if (blocks > sb.w4_resident_blocks) blocks = sb.w4_resident_blocks;
    const float* root_box = sb.root_box; QNode4* out = sb.qnodes4; int* queue = sb.w4_queue; unsigned* ctr = sb.w4_ctr;
    int mn = max_nodes, r2 = root2; bool fold = sb.fold_pairs && n > 2; QGrid* qgrid = sb.qgrid;
    void* args[] = {(void*)&walk, (void*)&r2, (void*)&mn, (void*)&root_box, (void*)&out, (void*)&queue, (void*)&ctr, (void*)&fold, (void*)&qgrid};
    e = cudaLaunchCooperativeKernel((const void*)wide_collapse_kernel, dim3(blocks), dim3(256), args, 0, stream);
    if (e != cudaSuccess) return (int)e;
    if (launches) *launches += 1;
    return (int)cudaGetLastError();
}

}  // namespace mcpt
