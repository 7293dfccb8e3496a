// lbvh.cu — seam 1 on the GPU: Morton codes -> radix sort -> Karras radix tree -> refit + layout.
// Replaces getMortonCode (MTPC/morton code.cpp:3-32), the std::sort by code (MTPC/MTPC.cpp:44,
// MTPC/sceneManagement.cpp:311-314) and the balanced-tree BVH (MTPC/BVH.cpp:44-132).
// This file is compiled WITHOUT fused multiply-add (--fmad=false, see build.py): K1's float pipeline
// must be IEEE-identical to the host's.
#include <cub/device/device_radix_sort.cuh>
#include "lbvh.h"
#include "quant.cuh"

namespace mcpt {

// ---- K1: Morton codes --------------------------------------------------------------------------
__device__ __forceinline__ unsigned expand_bits_10(unsigned v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
// (x - MINP) / (MAXP - MINP) with the integer macros MINP=-1, MAXP=4 (MTPC/morton code.h:6-7):
// float subtract, float divide by 5, then *1024, clamp to [0,1023], truncate.
__device__ __forceinline__ unsigned quantize_axis(float x) {
    float m = __fdiv_rn(__fsub_rn(x, -1.0f), 5.0f);
    m = __fmul_rn(m, 1024.0f);
    m = fminf(fmaxf(m, 0.0f), 1023.0f);
    return (unsigned)m;
}
// One thread per face.  The centroid is formed in FP64 from the FP64 vertices and only then narrowed
// to FP32, exactly like the call at MTPC/sceneManagement.cpp:176-179.
// Also zeroes the refit's arrival counters, and raises *bad when a coordinate is not finite or beyond +-1e18 (the build's FP32 box
// areas would overflow or turn NaN; mcpt_build_bvh then fails with MCPT_ERR_ARG instead of building a tree nothing can hit).
__global__ void __launch_bounds__(256) morton_kernel(const double* __restrict__ pos, int n, uint32_t* __restrict__ codes,
                                                     int* __restrict__ iota, unsigned* __restrict__ flags, unsigned* __restrict__ bad) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i < n - 1) flags[i] = 0u;
    const double* p = pos + 9 * (size_t)i;
    bool ok = true;
#pragma unroll
    for (int a = 0; a < 9; a++) ok = ok && (fabs(p[a]) <= 1.0e18);   // false for NaN and inf
    if (!ok) *bad = 1u;
    const double cx = __ddiv_rn(__dadd_rn(__dadd_rn(p[0], p[3]), p[6]), 3.0);
    const double cy = __ddiv_rn(__dadd_rn(__dadd_rn(p[1], p[4]), p[7]), 3.0);
    const double cz = __ddiv_rn(__dadd_rn(__dadd_rn(p[2], p[5]), p[8]), 3.0);
    const unsigned xx = expand_bits_10(quantize_axis(__double2float_rn(cx)));
    const unsigned yy = expand_bits_10(quantize_axis(__double2float_rn(cy)));
    const unsigned zz = expand_bits_10(quantize_axis(__double2float_rn(cz)));
    codes[i] = xx * 4u + yy * 2u + zz;  // x is the most significant axis (morton code.cpp:20)
    iota[i] = i;
}

// ---- K2: Karras 2012 radix tree over keys (code << 32 | sorted position) -------------------------
__device__ __forceinline__ int delta(const uint32_t* __restrict__ codes, int n, int i, unsigned long long ki, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long kj = ((unsigned long long)codes[j] << 32) | (unsigned)j;
    return __clzll((long long)(ki ^ kj));
}
__global__ void __launch_bounds__(256) radix_tree_kernel(const uint32_t* __restrict__ codes, int n, BvhNode* nodes,
                                                         int* __restrict__ leaf_parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const unsigned long long ki = ((unsigned long long)codes[i] << 32) | (unsigned)i;
    const int d = (delta(codes, n, i, ki, i + 1) - delta(codes, n, i, ki, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(codes, n, i, ki, i - d);
    int lmax = 2;
    while (delta(codes, n, i, ki, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(codes, n, i, ki, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(codes, n, i, ki, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(codes, n, i, ki, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int c0 = (lo == gamma) ? ~gamma : gamma;
    const int c1 = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    int* self = reinterpret_cast<int*>(&nodes[i].q[3]);
    self[0] = c0; self[1] = c1; self[3] = 0;
    if (i == 0) self[2] = -1;
    if (c0 < 0) leaf_parent[gamma] = i; else reinterpret_cast<int*>(&nodes[c0].q[3])[2] = i;
    if (c1 < 0) leaf_parent[gamma + 1] = i; else reinterpret_cast<int*>(&nodes[c1].q[3])[2] = i;
}

// ---- K3: leaf emit (traversal layout, K3a) and bottom-up AABB refit with atomic arrival counters (K3b) ----------
struct Box3 { float lo[3], hi[3]; };

__device__ __forceinline__ void store_child_box(BvhNode* node, int slot, const Box3& b) {
    // q[slot] = (lo.x, hi.x, lo.y, hi.y); q[2].xy or .zw = (lo.z, hi.z).  .cg stores/loads: L2 is the
    // point of coherence between the two sibling threads (they may sit on different SMs).
    __stcg(&node->q[slot], make_float4(b.lo[0], b.hi[0], b.lo[1], b.hi[1]));
    __stcg(reinterpret_cast<float2*>(&node->q[2]) + slot, make_float2(b.lo[2], b.hi[2]));
}
__device__ __forceinline__ Box3 load_child_box(const BvhNode* node, int slot) {
    const float4 xy = __ldcg(&node->q[slot]);
    const float2 z = __ldcg(reinterpret_cast<const float2*>(&node->q[2]) + slot);
    Box3 b;
    b.lo[0] = xy.x; b.hi[0] = xy.y; b.lo[1] = xy.z; b.hi[1] = xy.w; b.lo[2] = z.x; b.hi[2] = z.y;
    return b;
}

// FP32 box of face f's three vertices, pushed outward (> 3 ulp: covers the v0 + e1 rounding of the traversal triangle and the
// slab test's rounding).  Shared by K3a (leaf clusters of the quality pass) and K3b (refit), which therefore agree bit for bit.
__device__ __forceinline__ Box3 leaf_box(const float v[9]) {
    Box3 b;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float lo = fminf(fminf(v[a], v[3 + a]), v[6 + a]), hi = fmaxf(fmaxf(v[a], v[3 + a]), v[6 + a]);
        const float pad = 4.0e-7f * fmaxf(fabsf(lo), fabsf(hi)) + 1.0e-30f;
        b.lo[a] = lo - pad; b.hi[a] = hi + pad;
    }
    return b;
}

// K3a: one thread per sorted leaf: gathers the face's FP64 inputs once, writes the 48-B triangle, the 64-B shading record and
// the leaf's cluster box for the quality pass; also fills the collapse's queue with W4_EMPTY (saves a memset launch).
__global__ void __launch_bounds__(256) emit_kernel(int n, const double* __restrict__ pos, const double* __restrict__ nrm,
                                                   const double* __restrict__ uv, const int* __restrict__ mat_id,
                                                   const int* __restrict__ leaf_face, TriGeom* __restrict__ tri,
                                                   TriShade* __restrict__ shade, float4* __restrict__ cl_lo, float4* __restrict__ cl_hi,
                                                   int* __restrict__ w4_queue) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    w4_queue[k] = -1;
    const int f = leaf_face[k];
    const double* p = pos + 9 * (size_t)f;
    float v[9];
#pragma unroll
    for (int a = 0; a < 9; a++) v[a] = __double2float_rn(p[a]);
    TriGeom g;
    g.v0m = make_float4(v[0], v[1], v[2], __int_as_float(mat_id[f]));
    g.e1f = make_float4(v[3] - v[0], v[4] - v[1], v[5] - v[2], __int_as_float(f));
    g.e2 = make_float4(v[6] - v[0], v[7] - v[1], v[8] - v[2], 0.0f);
    tri[k] = g;
    const double* q = nrm + 9 * (size_t)f;
    const double* t = uv + 6 * (size_t)f;
    TriShade s;
    s.a = make_float4((float)q[0], (float)q[1], (float)q[2], (float)q[3]);
    s.b = make_float4((float)q[4], (float)q[5], (float)q[6], (float)q[7]);
    s.c = make_float4((float)q[8], (float)t[0], (float)t[1], (float)t[2]);
    s.d = make_float4((float)t[3], (float)t[4], (float)t[5], 0.0f);
    shade[k] = s;
    const Box3 b = leaf_box(v);
    cl_lo[k] = make_float4(b.lo[0], b.lo[1], b.lo[2], __int_as_float(~k));   // leaf cluster of the quality pass (ploc.cu)
    cl_hi[k] = make_float4(b.hi[0], b.hi[1], b.hi[2], 0.0f);
}

// K3b: bottom-up AABB refit of the radix tree with atomic arrival counters.  One thread per sorted leaf forms the leaf box and
// carries it up: at each parent the first child to arrive stores its box in the parent's slot and stops; the second one reads
// the sibling's box back (L2) and continues with the union.  Runs on the build's SIDE stream next to the quality pass, which
// needs only K3a's leaf boxes (lbvh_build); write_root: this tree is the one traversal walks, so its root box is the scene box.
__global__ void __launch_bounds__(256) refit_kernel(int n, const double* __restrict__ pos, const int* __restrict__ leaf_face,
                                                    const int* __restrict__ leaf_parent, BvhNode* nodes, unsigned* flags,
                                                    float* __restrict__ root_box, const bool write_root) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const double* p = pos + 9 * (size_t)leaf_face[k];
    float v[9];
#pragma unroll
    for (int a = 0; a < 9; a++) v[a] = __double2float_rn(p[a]);
    Box3 b = leaf_box(v);
    if (n == 1) {  // degenerate tree: node 0 points at the only leaf twice
        store_child_box(&nodes[0], 0, b); store_child_box(&nodes[0], 1, b);
        int* c = reinterpret_cast<int*>(&nodes[0].q[3]);
        c[0] = ~0; c[1] = ~0; c[2] = -1; c[3] = 0;
        for (int a = 0; a < 3; a++) { root_box[a] = b.lo[a]; root_box[3 + a] = b.hi[a]; }
        return;
    }
    int child = ~k;
    int parent = leaf_parent[k];
    while (true) {
        BvhNode* node = &nodes[parent];
        const int* c = reinterpret_cast<const int*>(&node->q[3]);
        const int slot = (c[0] == child) ? 0 : 1;
        store_child_box(node, slot, b);
        // release our box / acquire the sibling's with ONE acq_rel atomic at GPU scope instead of __threadfence() (a
        // sequentially-consistent MEMBAR.SC, the kernel's top stall in round 1's ncu capture) + a relaxed atomic
        unsigned arrived;
        asm volatile("atom.add.acq_rel.gpu.u32 %0, [%1], 1;" : "=r"(arrived) : "l"(&flags[parent]) : "memory");
        if (arrived == 0u) return;  // sibling subtree not finished yet
        const Box3 o = load_child_box(node, slot ^ 1);
#pragma unroll
        for (int a = 0; a < 3; a++) { b.lo[a] = fminf(b.lo[a], o.lo[a]); b.hi[a] = fmaxf(b.hi[a], o.hi[a]); }
        child = parent;
        parent = c[2];
        if (parent < 0) break;
    }
    if (write_root) for (int a = 0; a < 3; a++) { root_box[a] = b.lo[a]; root_box[3 + a] = b.hi[a]; }
}

// ---- K4: quantise the refitted nodes onto a 16-bit grid over the root box ----------------------------------
// One thread per internal node: 64-B FP32 node in, 32-B node out (grid and rounding rules: quant.cuh).
__global__ void __launch_bounds__(256) quantize_nodes_kernel(int n_nodes, const BvhNode* __restrict__ nodes,
                                                             const float* __restrict__ root_box, QNode* __restrict__ qnodes,
                                                             QGrid* __restrict__ qgrid, const bool fold) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    float o[3], cell[3], inv[3];
    quant_grid(root_box, o, cell, inv);
    if (i == 0) {
        qgrid->origin = make_float4(o[0], o[1], o[2], 0.0f);
        qgrid->cell = make_float4(cell[0], cell[1], cell[2], 0.0f);
    }
    const float4 q0 = nodes[i].q[0], q1 = nodes[i].q[1], q2 = nodes[i].q[2], q3 = nodes[i].q[3];
    QNode out;
    out.c[0] = make_uint4(quant_lo(q0.x, o[0], inv[0]) | (quant_hi(q0.y, o[0], inv[0]) << 16),
                          quant_lo(q0.z, o[1], inv[1]) | (quant_hi(q0.w, o[1], inv[1]) << 16),
                          quant_lo(q2.x, o[2], inv[2]) | (quant_hi(q2.y, o[2], inv[2]) << 16), (unsigned)fold_leaf_pair(nodes, __float_as_int(q3.x), fold));
    out.c[1] = make_uint4(quant_lo(q1.x, o[0], inv[0]) | (quant_hi(q1.y, o[0], inv[0]) << 16),
                          quant_lo(q1.z, o[1], inv[1]) | (quant_hi(q1.w, o[1], inv[1]) << 16),
                          quant_lo(q2.z, o[2], inv[2]) | (quant_hi(q2.w, o[2], inv[2]) << 16), (unsigned)fold_leaf_pair(nodes, __float_as_int(q3.y), fold));
    qnodes[i] = out;
}

// ---- sort ------------------------------------------------------------------------------------------------------------
// cub::DeviceRadixSort::SortPairs over 30 key bits (a single-tile kernel below ~4.8 k keys, histogram + scan + four onesweep
// passes above): LSD, hence stable — ties keep face order, the canonical leaf order.  Measured and dropped: one CTA of 1024
// threads x 16 keys with cub::BlockRadixSort for up to 16 Ki keys (70.8 us for cornell-box's 15 056 keys against 62 us here); an
// all-pairs rank sort for n <= 16 Ki (rank = #{j : code_j < code_i + (j < i)}, segments in shared memory, two launches): correct,
// 51 + 4 us against 57 us — 2.3e8 compare-and-count pairs are ~3 half-rate ALU instructions each.
static int sort_codes(SceneBuffers& sb, cudaStream_t stream, int* nl) {
    const int n = sb.n_faces;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(sb.cub_temp, sb.cub_temp_bytes, sb.codes_obj, sb.codes_sorted, sb.iota,
                                                    sb.leaf_face, n, 0, 30, stream);
    *nl += n <= 4864 ? 1 : 6;
    return (int)e;
}

// ---- host-side orchestration -----------------------------------------------------------------------
size_t lbvh_sort_temp_bytes(int n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const int*)nullptr,
                                    (int*)nullptr, n, 0, 30);
    return bytes;
}

// Two streams: the radix tree (K2) and its refit (K3b) need only the sorted codes; the quality pass needs only K3a's leaf boxes.
// So after the sort the build forks: SIDE: K2 -> K3b;  MAIN: K3a -> PLOC rounds -> finish -> wide collapse; and joins at the
// end (cornell-box: ~55 us of the radix tree's latency chain hidden behind ~150 us of clustering + collapse).
int lbvh_build(SceneBuffers& sb, cudaStream_t stream, const BuildStreams& bs, int* launches) {
    const int n = sb.n_faces;
    const int tpb = 256;
    int nl = 0;
    cudaError_t e = cudaMemsetAsync(sb.w4_ctr, 0, 16 * sizeof(unsigned), stream);   // [5] = K1's bad-vertex flag
    if (e != cudaSuccess) return (int)e;
    morton_kernel<<<(n + tpb - 1) / tpb, tpb, 0, stream>>>(sb.pos, n, sb.codes_obj, sb.iota, sb.flags, sb.w4_ctr + 5); nl++;
    // stable LSD radix sort of (code, face index), 30 key bits: ties keep face-index order
    { const int se = sort_codes(sb, stream, &nl); if (se) return se; }
    const bool ploc = sb.use_ploc && n >= 2;
    if ((e = cudaEventRecord(bs.ev_fork, stream)) != cudaSuccess) return (int)e;
    if ((e = cudaStreamWaitEvent(bs.side, bs.ev_fork, 0)) != cudaSuccess) return (int)e;
    if (n > 1) { radix_tree_kernel<<<(n - 1 + tpb - 1) / tpb, tpb, 0, bs.side>>>(sb.codes_sorted, n, sb.nodes, sb.leaf_parent); nl++; }
    refit_kernel<<<(n + tpb - 1) / tpb, tpb, 0, bs.side>>>(n, sb.pos, sb.leaf_face, sb.leaf_parent, sb.nodes, sb.flags, sb.root_box, !ploc); nl++;
    if ((e = cudaEventRecord(bs.ev_join, bs.side)) != cudaSuccess) return (int)e;
    emit_kernel<<<(n + tpb - 1) / tpb, tpb, 0, stream>>>(n, sb.pos, sb.nrm, sb.uv, sb.mat_id, sb.leaf_face, sb.tri, sb.shade,
                                                         sb.cl_lo[0], sb.cl_hi[0], sb.w4_queue); nl++;
    const BvhNode* walk = sb.nodes;
    sb.root = 0; sb.walk_is_radix = true;
    bool w4_prepared = false;
    if (ploc) {
        const int pe = ploc_build(sb, stream, &nl, MCPT_WIDE != 0);
        if (pe) return pe;
        walk = sb.nodes2; sb.root = n - 2; sb.walk_is_radix = false;
        w4_prepared = MCPT_WIDE != 0;
    } else {
        if ((e = cudaStreamWaitEvent(stream, bs.ev_join, 0)) != cudaSuccess) return (int)e;   // the collapse walks the radix tree
    }
#if MCPT_WIDE
    { const int we = wide_build(sb, walk, sb.root, stream, &nl, w4_prepared); if (we) return we; }
#else
    (void)w4_prepared;
    const int n_nodes = n > 1 ? n - 1 : 1;
    quantize_nodes_kernel<<<(n_nodes + tpb - 1) / tpb, tpb, 0, stream>>>(n_nodes, walk, sb.root_box, sb.qnodes, sb.qgrid, sb.fold_pairs && n > 2); nl++;
#endif
    if (ploc && (e = cudaStreamWaitEvent(stream, bs.ev_join, 0)) != cudaSuccess) return (int)e;      // join
    if (launches) *launches = nl;
    sb.built = true;
    return (int)cudaGetLastError();
}

// The radix tree as the traversal tree (the quality pass stalled, or its tree is too deep for the traversal stack).  K3 left
// the scene box in root_box (the radix root's box; the PLOC finish writes the same union), so only the collapse is redone.
int lbvh_fallback_to_radix(SceneBuffers& sb, cudaStream_t stream, int* launches) {
    const int n = sb.n_faces;
    int nl = 0;
    sb.root = 0; sb.walk_is_radix = true;
#if MCPT_WIDE
    { const int we = wide_build(sb, sb.nodes, 0, stream, &nl, false); if (we) return we; }
#else
    const int n_nodes = n > 1 ? n - 1 : 1;
    quantize_nodes_kernel<<<(n_nodes + 255) / 256, 256, 0, stream>>>(n_nodes, sb.nodes, sb.root_box, sb.qnodes, sb.qgrid, sb.fold_pairs && n > 2); nl++;
#endif
    (void)n;
    if (launches) *launches += nl;
    return (int)cudaGetLastError();
}

}  // namespace mcpt
