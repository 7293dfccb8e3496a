// Device-side data model shared by the build (lbvh.cu), traversal/integrator (render.cu) and the
// C ABI (api.cu).  Everything the kernels touch lives in HBM in the layouts described in DESIGN.md.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/mcpt.h"

namespace mcpt {

// BVH node, 64 B = 4 x float4 (one 128-B line holds two nodes; every field is a 16-B vector load):
//   q[0] = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)
//   q[1] = (c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y)
//   q[2] = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)
//   q[3] = (child0, child1, parent, unused) as int bits; child >= 0: internal node, child < 0: leaf ~child
struct __align__(16) BvhNode { float4 q[4]; };

// Traversal node, 32 B = 2 x uint4 (ONE 32-B sector per visit): the two child boxes on a 16-bit grid over the root box,
//   c[s] = (lo.x | hi.x << 16, lo.y | hi.y << 16, lo.z | hi.z << 16, child index as in BvhNode)
// A child that is a bottom node over two single triangles k < k' with k' - k <= 15 is folded into this node as the leaf
// reference ~(k | (k' - k) << 26): the two triangles of a quad have almost the same box, so a ray that enters one nearly
// always enters the other, and one reference saves the node visit and the second stop of the lane (MCPT_LEAF_BITS = 26).
// lo rounded down and hi rounded up by two cells, so the quantised box contains the FP32 box with room for the
// decode's rounding (DESIGN.md "quantised nodes").  Half the loads of BvhNode: traversal is bound by L1 wavefronts.
struct __align__(16) QNode { uint4 c[2]; };
#define MCPT_LEAF_BITS 26
// 4-wide traversal node, 64 B = 4 x uint4, same per-child encoding as QNode; an unused slot holds an inverted box (never
// hit).  Built from the binary tree by K4b (wide.cu), breadth-first, root = node 0.  MCPT_WIDE selects which tree the
// traversal kernels walk (compile time: the node step is the innermost loop).
struct __align__(16) QNode4 { uint4 c[4]; };
#ifndef MCPT_WIDE
#define MCPT_WIDE 1
#endif
// Grid of the quantisation, 8 floats: origin.xyz, 0, cell.xyz, 0  (world = origin + q * cell)
struct __align__(16) QGrid { float4 origin, cell; };

// Triangle, 48 B = 3 x float4, Morton-sorted leaf order: (v0, mat_id), (e1 = v1-v0, face_id), (e2 = v2-v0, 0)
struct __align__(16) TriGeom { float4 v0m, e1f, e2; };

// Shading attributes, 64 B = 4 x float4, same order:
//   a = (n0.xyz, n1.x)  b = (n1.yz, n2.xy)  c = (n2.z, uv0.xy, uv1.x)  d = (uv1.y, uv2.xy, 0)
struct __align__(16) TriShade { float4 a, b, c, d; };

// Light-sampling record per light triangle in OBJ (= CDF) order, 80 B: positions and vertex normals.
struct __align__(16) LightTri { float4 p0n0x, p1n0y, p2n0z, n1n2x, n2yz; };

struct DMaterial {
    float kd[3]; float kd_norm;
    float ks[3]; float ks_norm;
    float Ns, Ni;
    int tex_id;   // < 0: none
    int light_id; // < 0: not emissive (light_map lookup, MTPC/pathTracing.cpp:141)
};
struct DTexture { int height, width; unsigned long long offset; };  // offset into tex_bytes
struct DLight {
    float radiance[3];
    float area;       // total area (pdf_light = 1/area, MTPC/pathTracing.cpp:218)
    int mat_id;
    int first;        // first entry in light_tris / light_cdf
    int count;        // triangles of that material
    int n_search;     // entries whose CDF can be selected by rnd in [0, area of light 0) (quirk Q6)
    float box_lo[3], box_hi[3];   // AABB of every triangle carrying this light's material, padded outward (shadow-ray early-out)
    int _pad[2];
};

struct DeviceScene {
    int n_faces;
    int n_materials, n_textures, n_lights;
    int root;                 // >= 0 internal node, < 0 leaf (single-triangle scene)
    int n_wide, wide_levels;  // wide nodes in use; internal levels of the wide tree (<= MCPT_MAX_WIDE_LEVELS)
    float area0;              // area of light 0: the frozen range of the selection uniform (quirk Q6)
    const BvhNode* nodes;     // FP32 nodes (build product, structural tests)
    const QNode* qnodes;      // 16-bit quantised nodes (what traversal reads)
    const QNode4* qnodes4;    // 4-wide quantised nodes (MCPT_WIDE)
    const QGrid* qgrid;
    const TriGeom* tri;
    const TriShade* shade;
    const int* leaf_face;     // sorted leaf -> OBJ face index
    const double* pos;        // FP64 vertex positions, OBJ order (9 per face): final-hit refinement of seam 2
    const DMaterial* materials;
    const DTexture* textures;
    const unsigned char* tex_bytes;
    const DLight* lights;
    const LightTri* light_tris;
    const float* light_cdf;
};

// Pinhole camera of generateImg (MTPC/pathTracing.cpp:276-294), basis precomputed on the host in double.
struct DCamera {
    double eye[3], start[3], pdx[3], pdy[3];  // pixel (j,i) point = start - pdy*i + pdx*j
    int width, height;
};

// n / d for 32-bit unsigned n and a divisor known on the host (Granlund-Montgomery round-up multiplier, branch-free form):
// four integer instructions instead of the ~20 of a division by a kernel parameter (wf_shade maps slot -> tile -> pixel with four).
struct FastDiv { unsigned mul, shift, d; };
inline FastDiv make_fastdiv(unsigned d) {
    FastDiv f; f.d = d;
    if (d <= 1u) { f.mul = 0u; f.shift = 0u; return f; }   // handled by the d == 1 branch of fastdiv
    unsigned s = 0; while ((1ull << s) < d) s++;             // ceil(log2 d)
    f.mul = (unsigned)((((1ull << s) - d) << 32) / d + 1ull);
    f.shift = s - 1u;
    return f;
}
__host__ __device__ __forceinline__ unsigned fastdiv(unsigned n, const FastDiv f) {
    if (f.d <= 1u) return n;
#ifdef __CUDA_ARCH__
    const unsigned q = __umulhi(f.mul, n);
#else
    const unsigned q = (unsigned)(((unsigned long long)f.mul * n) >> 32);
#endif
    return (((n - q) >> 1) + q) >> f.shift;
}

struct RenderParams {
    DeviceScene scene;
    DCamera cam;
    int spp;            // total samples per pixel of the whole (multi-GPU) render
    int chunk_begin;    // first sample chunk rendered by this launch
    int n_chunks;       // chunks rendered by this launch
    int chunks_per_thread;
    FastDiv div_tiles, div_tiles_x;   // by the number of 8x4 pixel tiles, by the tiles per row
    int flags;                        // MCPT_FLAG_* of the render (the quality extensions are run-time switches)
    unsigned seed_lo, seed_hi;
    const float* cam_dirs; // height*width*3: primary ray directions from raygen_kernel (K7)
    float* planes;      // n_chunks planes of height*width*3 floats: per-chunk SUM of sample radiance
    unsigned long long* counters; // [0] rays, [1] node visits, [2] triangle tests
};

}  // namespace mcpt
