// quant.cuh — the 16-bit node grid (QGrid) and leaf-reference folding shared by K4 (lbvh.cu) and K4b (wide.cu).
#pragma once
#include "device_types.cuh"

namespace mcpt {

// lo is floored and hi is ceiled, then both are pushed out by two more cells: one for the float rounding of
// (x - origin) / cell here, one for the half-cell error of the magic-number decode in node_step (trace.cuh).
__device__ __forceinline__ unsigned quant_lo(float x, float o, float inv_cell) {
    const float q = floorf((x - o) * inv_cell) - 2.0f;
    return (unsigned)fminf(fmaxf(q, 0.0f), 65535.0f);
}
__device__ __forceinline__ unsigned quant_hi(float x, float o, float inv_cell) {
    const float q = ceilf((x - o) * inv_cell) + 2.0f;
    return (unsigned)fminf(fmaxf(q, 0.0f), 65535.0f);
}
// Grid over the root box: it spans cells 8 .. 65528, leaving room for the outward push.
__device__ __forceinline__ void quant_grid(const float* __restrict__ root_box, float o[3], float cell[3], float inv[3]) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float lo = root_box[a], hi = root_box[3 + a];
        const float ext = fmaxf(hi - lo, 1.0e-20f);
        cell[a] = ext * (1.0f / 65520.0f);
        o[a] = lo - 8.0f * cell[a];
        inv[a] = 1.0f / cell[a];
    }
}
// Child reference of a traversal node: a bottom node over two single triangles becomes one two-triangle leaf reference.
__device__ __forceinline__ int fold_leaf_pair(const BvhNode* __restrict__ nodes, const int c, const bool fold) {
    if (c < 0 || !fold) return c;
    const float4 q3 = nodes[c].q[3];
    const int a = __float_as_int(q3.x), b = __float_as_int(q3.y);
    if (a >= 0 || b >= 0) return c;
    const int ka = ~a, kb = ~b, lo = min(ka, kb), d = max(ka, kb) - lo;
    if (d == 0 || d > 15 || lo >= (1 << MCPT_LEAF_BITS)) return c;
    return (int)~((unsigned)lo | ((unsigned)d << MCPT_LEAF_BITS));
}

}  // namespace mcpt
