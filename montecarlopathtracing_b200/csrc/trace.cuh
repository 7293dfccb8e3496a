// trace.cuh — seam 2 on the GPU: closest-hit BVH traversal + Möller–Trumbore.
// Replaces bvh_intersect / ray_intersect (MTPC/pathTracing.cpp:334-390), the slab test
// intersect(Ray&, boundingBox&) (MTPC/sceneManagement.cpp:340-391), the plane + same-side triangle
// test intersect(Ray&, Face&, Vertex&) (:316-338) and findGarCor (MTPC/pathTracing.cpp:394-432, whose
// (b2,b3) are Möller–Trumbore's (u,v)).
//
// Semantics kept from the reference: two-sided triangles, inclusive edges, hits need t > 0 (no
// epsilon), boxes are clipped to t >= 0, the closest hit wins and an exact tie in t goes to the
// lowest sorted-leaf index (first found in the reference's left-to-right DFS, :355-361), and a ray
// whose d.x*t vanishes against o.x in FP64 hits nothing (its t = (p.x - o.x)/d.x is 0 or NaN, :347 — quirk Q2).
// New: children are visited near-first and pruned against the current best t.
#pragma once
#include "device_types.cuh"
#include "vecmath.cuh"

namespace mcpt {

struct Hit { float t, u, v; int k; };  // k = sorted-leaf index, -1 = miss

#define MCPT_STACK_SIZE 64   // arrays are declared with MCPT_STACK_SIZE + 4 entries (node4_step's branch-free pushes)

// Edge tolerance of the FP32 barycentric test.  The reference's FP64 same-side test is inclusive (>= 0), so a ray
// through a shared edge hits one of the two triangles; in FP32 the rounding of u, v can put it just outside BOTH
// (a crack: the image diagonals of cornell-box run exactly along wall edges).  Accepting |excess| <= 2e-6 closes the
// cracks; which of two edge-sharing triangles wins is then decided by t like in the reference.
#define MCPT_EDGE_EPS 2.0e-6f

__device__ __forceinline__ bool a_pos(float a) { return !(a < 0.0f); }
// MUFU.RCP alone (1 ulp) instead of the IEEE reciprocal's 8-instruction sequence; both integrators and seam 2 share it.
__device__ __forceinline__ float __frcp_approx(const float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// Per-ray constants.  Slab planes come from the 16-bit grid of the quantised nodes: a plane with grid coordinate q is
// crossed at t = (origin + q*cell - o) / d = q*a + b with a = cell/d, b = (origin - o)/d.  q is turned into a float by
// the magic-number trick — PRMT drops its 16 bits into the mantissa of 2^23, giving the float 2^23 + q exactly — so
// t = (2^23 + q)*a + (b - 2^23*a): one PRMT and one FMA per plane.  b' = b - 2^23*a is rounded to half a grid cell
// at worst, which the two-cell outward push of the quantiser covers (lbvh.cu K4).
struct RayCtx { float3 o, d; float ax, ay, az, bx, by, bz; unsigned sx, sy, sz; };
// sm_100a issues packed pairs of FP32 FMAs as ONE instruction (FFMA2, PTX fma.rn.f32x2): the slab planes of a child pair up as
// (near x, near y), (far x, far y), (near z, far z): three FFMA2 instead of six FFMA per child.  Measured and NOT kept
// (-DMCPT_FFMA2=1 builds it): the operands must sit in aligned 64-bit register pairs, wf_trace starts spilling at its 56-register
// budget, and throughput DROPS: cornell-box 1 805 -> 1 738 Msamples/s, veach-mis 736 -> 717, 5 M triangles 974 -> 915.
#ifndef MCPT_FFMA2
#define MCPT_FFMA2 0
#endif
__device__ __forceinline__ unsigned long long pack2(const float lo, const float hi) {
    unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r;
}
__device__ __forceinline__ void fma2(const float a0, const float a1, const unsigned long long b, const unsigned long long c, float& r0, float& r1) {
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pack2(a0, a1)), "l"(b), "l"(c));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r0), "=f"(r1) : "l"(r));
}

__device__ __forceinline__ RayCtx ray_setup(const float3 o, const float3 d, const QGrid* __restrict__ grid) {
    RayCtx r; r.o = o; r.d = d;
    const float4 go = __ldg(&grid->origin), gc = __ldg(&grid->cell);
    const float ooeps = 1.0e-24f;  // avoid inf/NaN slabs for axis-parallel rays
    // MUFU.RCP alone: a 1-ulp error in 1/d scales every slab distance by 1 +- 1.2e-7, far inside the two-cell push of the boxes
    const float idx = __frcp_approx(fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x));
    const float idy = __frcp_approx(fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y));
    const float idz = __frcp_approx(fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z));
    r.ax = gc.x * idx; r.ay = gc.y * idy; r.az = gc.z * idz;
    r.bx = fmaf(-8388608.0f, r.ax, (go.x - o.x) * idx);
    r.by = fmaf(-8388608.0f, r.ay, (go.y - o.y) * idy);
    r.bz = fmaf(-8388608.0f, r.az, (go.z - o.z) * idz);
    // PRMT selectors of the NEAR plane per axis: the low half (lo) when the ray travels in +axis, else the high half
    // (hi); the far plane's selector is the near one ^ 0x0022.  Saves the six min/max per child that order lo and hi.
    r.sx = a_pos(r.ax) ? 0x7410u : 0x7432u; r.sy = a_pos(r.ay) ? 0x7410u : 0x7432u; r.sz = a_pos(r.az) ? 0x7410u : 0x7432u;
    return r;
}

// Stack bottom: negative like a leaf, never a valid ~k.
#define MCPT_SENT ((int)0x80000000)

__device__ __forceinline__ float q_lo(unsigned w) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7410)); }
__device__ __forceinline__ float q_hi(unsigned w) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7432)); }
__device__ __forceinline__ float q_sel(unsigned w, unsigned sel) { return __uint_as_float(__byte_perm(w, 0x4B000000u, sel)); }

// One internal-node step: fetch the 32-B node (two 16-B loads, one sector), slab-test both children against
// [0, best_t], descend into the nearer hit child (pushing the farther) or pop.  `node` >= 0 on entry.
template <bool COUNT>
__device__ __forceinline__ void node_step(const QNode* __restrict__ nodes, int& node, const RayCtx& r, const float best_t,
                                          int* stack, int& sp, unsigned& n_nodes) {
    const uint4 k0 = __ldg(&nodes[node].c[0]);
    const uint4 k1 = __ldg(&nodes[node].c[1]);
    if (COUNT) n_nodes++;
    const unsigned fx = r.sx ^ 0x0022u, fy = r.sy ^ 0x0022u, fz = r.sz ^ 0x0022u;
    const float n0x = fmaf(q_sel(k0.x, r.sx), r.ax, r.bx), f0x = fmaf(q_sel(k0.x, fx), r.ax, r.bx);
    const float n0y = fmaf(q_sel(k0.y, r.sy), r.ay, r.by), f0y = fmaf(q_sel(k0.y, fy), r.ay, r.by);
    const float n0z = fmaf(q_sel(k0.z, r.sz), r.az, r.bz), f0z = fmaf(q_sel(k0.z, fz), r.az, r.bz);
    const float c0min = fmaxf(fmaxf(n0x, n0y), fmaxf(n0z, 0.0f));
    const float c0max = fminf(fminf(f0x, f0y), fminf(f0z, best_t));
    const float n1x = fmaf(q_sel(k1.x, r.sx), r.ax, r.bx), f1x = fmaf(q_sel(k1.x, fx), r.ax, r.bx);
    const float n1y = fmaf(q_sel(k1.y, r.sy), r.ay, r.by), f1y = fmaf(q_sel(k1.y, fy), r.ay, r.by);
    const float n1z = fmaf(q_sel(k1.z, r.sz), r.az, r.bz), f1z = fmaf(q_sel(k1.z, fz), r.az, r.bz);
    const float c1min = fmaxf(fmaxf(n1x, n1y), fmaxf(n1z, 0.0f));
    const float c1max = fminf(fminf(f1x, f1y), fminf(f1z, best_t));
    const bool h0 = c0min <= c0max, h1 = c1min <= c1max;
    const int c0 = (int)k0.w, c1 = (int)k1.w;
    if (h0 && h1) {
        const bool swap = c1min < c0min;   // near child first
        node = swap ? c1 : c0;
        if (sp < MCPT_STACK_SIZE) stack[sp++] = swap ? c0 : c1;
    } else if (h0 || h1) {
        node = h0 ? c0 : c1;
    } else {
        node = stack[--sp];
    }
}

// One 4-wide node step (QNode4): four 16-B loads, four independent slab tests, then the hit children in near-to-far order —
// the nearest is descended into, the others are pushed farthest first.  Order keys: the entry distance's float bits
// (>= 0, so unsigned order = float order); a miss is 0xFFFFFFFF and sorts last.  Keys and child references are sorted as
// PAIRS (one compare + four selects per exchange): cheaper than sorting slot-tagged keys and picking the reference of every
// sorted slot afterwards.  The pushes are predicated stores, not branches (ncu source view: the three branchy push blocks
// ran on nearly every step for 2-6 of 32 lanes, ~40 of the step's ~150 instructions).
__device__ __forceinline__ unsigned slab_key(const uint4 k, const RayCtx& r, const unsigned fx, const unsigned fy, const unsigned fz,
                                             const float best_t) {
#if MCPT_FFMA2
    float nx, ny, gx, gy, nz, gz;
    const unsigned long long axy = pack2(r.ax, r.ay), bxy = pack2(r.bx, r.by), azz = pack2(r.az, r.az), bzz = pack2(r.bz, r.bz);
    fma2(q_sel(k.x, r.sx), q_sel(k.y, r.sy), axy, bxy, nx, ny);
    fma2(q_sel(k.x, fx), q_sel(k.y, fy), axy, bxy, gx, gy);
    fma2(q_sel(k.z, r.sz), q_sel(k.z, fz), azz, bzz, nz, gz);
#else
    const float nx = fmaf(q_sel(k.x, r.sx), r.ax, r.bx), gx = fmaf(q_sel(k.x, fx), r.ax, r.bx);
    const float ny = fmaf(q_sel(k.y, r.sy), r.ay, r.by), gy = fmaf(q_sel(k.y, fy), r.ay, r.by);
    const float nz = fmaf(q_sel(k.z, r.sz), r.az, r.bz), gz = fmaf(q_sel(k.z, fz), r.az, r.bz);
#endif
    const float cmin = fmaxf(fmaxf(nx, ny), fmaxf(nz, 0.0f));
    const float cmax = fminf(fminf(gx, gy), fminf(gz, best_t));
    return cmin <= cmax ? __float_as_uint(cmin) : 0xFFFFFFFFu;
}
__device__ __forceinline__ void pair_cswap(unsigned& ka, int& ra, unsigned& kb, int& rb) {
    const bool p = kb < ka;
    const unsigned k = p ? kb : ka; kb = p ? ka : kb; ka = k;
    const int r = p ? rb : ra; rb = p ? ra : rb; ra = r;
}
// Where a traversal stack lives.  LocalStack: the thread's local-memory array (lanes at different depths touch different 128-B
// lines: one L1 wavefront per distinct depth in the warp).  SmemStack (wf_trace with -DMCPT_SMEM_STACK=1): entry i of thread t
// at word i * blockDim + t of the CTA's dynamic shared memory — bank = lane whatever the depth, one wavefront per access.
struct LocalStack {
    int* a;
    __device__ __forceinline__ int get(const int i) const { return a[i]; }
    __device__ __forceinline__ void put(const int i, const int v) const { a[i] = v; }
};
struct SmemStack {
    int* base; int stride;   // base = &smem[threadIdx.x], stride = blockDim.x
    __device__ __forceinline__ int get(const int i) const { return base[i * stride]; }
    __device__ __forceinline__ void put(const int i, const int v) const { base[i * stride] = v; }
};
// The top MCPT_SMEM_TOP wide nodes (breadth-first layout: the top levels are the first nodes) staged in shared memory, 16-B
// chunks rotated by the node index so that lanes reading chunk c of different nodes spread over the bank groups.
#ifndef MCPT_SMEM_TOP
#define MCPT_SMEM_TOP 0
#endif
#ifndef MCPT_SMEM_STACK
#define MCPT_SMEM_STACK 0
#endif
__device__ __forceinline__ int top_slot(const int node, const int c) { return node * 4 + ((c + (node >> 1)) & 3); }

template <bool COUNT, class Stack>
__device__ __forceinline__ void node4_step(const QNode4* __restrict__ nodes, int& node, const RayCtx& r, const float best_t,
                                           const Stack stack, int& sp, unsigned& n_nodes, const uint4* __restrict__ top = nullptr) {
    uint4 k0, k1, k2, k3;
    if (MCPT_SMEM_TOP && top && node < MCPT_SMEM_TOP) {
        k0 = top[top_slot(node, 0)]; k1 = top[top_slot(node, 1)]; k2 = top[top_slot(node, 2)]; k3 = top[top_slot(node, 3)];
    } else {
        k0 = __ldg(&nodes[node].c[0]); k1 = __ldg(&nodes[node].c[1]); k2 = __ldg(&nodes[node].c[2]); k3 = __ldg(&nodes[node].c[3]);
    }
    if (COUNT) n_nodes++;
    const unsigned fx = r.sx ^ 0x0022u, fy = r.sy ^ 0x0022u, fz = r.sz ^ 0x0022u;
    unsigned s0 = slab_key(k0, r, fx, fy, fz, best_t), s1 = slab_key(k1, r, fx, fy, fz, best_t);
    unsigned s2 = slab_key(k2, r, fx, fy, fz, best_t), s3 = slab_key(k3, r, fx, fy, fz, best_t);
    int r0 = (int)k0.w, r1 = (int)k1.w, r2 = (int)k2.w, r3 = (int)k3.w;
    pair_cswap(s0, r0, s1, r1); pair_cswap(s2, r2, s3, r3); pair_cswap(s0, r0, s2, r2); pair_cswap(s1, r1, s3, r3); pair_cswap(s1, r1, s2, r2);
    // no bounds test: mcpt_build_bvh rejects (or rebuilds) a wide tree with more than MCPT_MAX_WIDE_LEVELS = 21 internal levels, and
    // a depth-first walk holds at most the sentinel + 3 siblings per level above the current node: sp <= 1 + 3*20 before the pushes
    if (s3 != 0xFFFFFFFFu) { stack.put(sp, r3); sp++; }
    if (s2 != 0xFFFFFFFFu) { stack.put(sp, r2); sp++; }
    if (s1 != 0xFFFFFFFFu) { stack.put(sp, r1); sp++; }
    node = r0;
    if (s0 == 0xFFFFFFFFu) node = stack.get(--sp);
}

// Shadow rays (MTPC/pathTracing.cpp:202-215) ask one question: does the CLOSEST hit carry the light's material?  The answer is
// known to be "no" as soon as a triangle of ANOTHER material is hit in front of the box that holds every triangle of the light's
// material (DLight::box_*): nothing of that material can be closer.  occl_t = the ray's entry distance into that box, shrunk a
// little (so rounding cannot make it late); a non-shadow ray has occl_t < 0 and the test never fires.  Exact, not an any-hit
// approximation: the visibility bit — hence the image — is the same with and without it (tests: bit-identity of the integrators
// with MCPT_FLAG_NO_SHADOW_EARLY_OUT).
struct Occluder { float t; int light_mat; };
__device__ __forceinline__ Occluder no_occluder() { Occluder o; o.t = -1.0f; o.light_mat = -1; return o; }
__device__ __forceinline__ Occluder shadow_occluder(const DLight& L, const float3 o, const float3 d, const float tmax) {
    const float ooeps = 1.0e-24f;
    const float ix = __frcp_approx(fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x));
    const float iy = __frcp_approx(fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y));
    const float iz = __frcp_approx(fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z));
    const float x0 = (L.box_lo[0] - o.x) * ix, x1 = (L.box_hi[0] - o.x) * ix;
    const float y0 = (L.box_lo[1] - o.y) * iy, y1 = (L.box_hi[1] - o.y) * iy;
    const float z0 = (L.box_lo[2] - o.z) * iz, z1 = (L.box_hi[2] - o.z) * iz;
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
    Occluder oc; oc.light_mat = L.mat_id;
    // a ray that misses the box cannot hit the light's material at all: every hit occludes
    oc.t = (tn <= tf * 1.00001f + 1.0e-6f) ? tn * 0.99999f - 1.0e-6f : tmax;
    return oc;
}

// Möller–Trumbore on sorted triangle k; keeps it in `best` if it is the closer hit (ties: lower leaf index).
// Returns true when a shadow ray's question is settled by this hit (see Occluder).
template <bool COUNT>
__device__ __forceinline__ bool tri_test(const TriGeom* __restrict__ tris, const int k, const RayCtx& r, Hit& best, unsigned& n_tris,
                                         const Occluder oc = Occluder{-1.0f, -1}) {
    const float4 v0 = __ldg(&tris[k].v0m);
    const float4 e1 = __ldg(&tris[k].e1f);
    const float4 e2 = __ldg(&tris[k].e2);
    if (COUNT) n_tris++;
    const float3 E1 = f3(e1), E2 = f3(e2);
    const float3 pvec = cross(r.d, E2);
    const float inv_det = __frcp_approx(dot(E1, pvec));  // det == 0 -> inf/NaN -> every comparison below fails
    const float3 tvec = r.o - f3(v0);
    const float u = dot(tvec, pvec) * inv_det;
    const float3 qvec = cross(tvec, E1);
    const float v = dot(r.d, qvec) * inv_det;
    const float t = dot(E2, qvec) * inv_det;
    if (u >= -MCPT_EDGE_EPS && v >= -MCPT_EDGE_EPS && u + v <= 1.0f + MCPT_EDGE_EPS && t > 0.0f && (t < best.t || (t == best.t && k < best.k))) {
        // quirk Q2 in full: the reference keeps a hit only if t = (p.x - o.x)/d.x > 0 with p.x = o.x + d.x*t formed in
        // FP64 (MTPC/pathTracing.cpp:347,355); when |d.x*t| is below half an ulp of o.x that is 0 and the hit is
        // dropped.  Candidates that pass the FP32 test are rare (1-2 per ray), so the FP64 check is cheap.
        // FP32 shortcut: |d.x*t| > 1e-12 |o.x| is ~1e4 FP64 ulps of o.x, so the FP64 sum certainly differs; the FP64 test runs
        // only for the few candidates at grazing d.x.
        if (fabsf(r.d.x * t) > 1.0e-12f * fabsf(r.o.x) || __dadd_rn((double)r.o.x, __dmul_rn((double)r.d.x, (double)t)) != (double)r.o.x) {
            best.t = t; best.u = u; best.v = v; best.k = k;
            return t < oc.t && __float_as_int(v0.w) != oc.light_mat;
        }
    }
    return false;
}

// Leaf references of the traversal nodes (device_types.cuh QNode): ~(k | delta << 26) = triangle k and, when delta != 0,
// triangle k + delta (a bottom node whose two children were single triangles, folded into its parent by K4).
__device__ __forceinline__ int leaf_first(const int ref) { return ~ref & ((1 << MCPT_LEAF_BITS) - 1); }
__device__ __forceinline__ int leaf_delta(const int ref) { return (~ref >> MCPT_LEAF_BITS) & 15; }
template <bool COUNT>
__device__ __forceinline__ bool leaf_test(const TriGeom* __restrict__ tris, const int ref, const RayCtx& r, Hit& best, unsigned& n_tris,
                                          const Occluder oc = Occluder{-1.0f, -1}) {
    const int k = leaf_first(ref), dk = leaf_delta(ref);
    bool settled = tri_test<COUNT>(tris, k, r, best, n_tris, oc);
    if (dk) settled = tri_test<COUNT>(tris, k + dk, r, best, n_tris, oc) || settled;
    return settled;
}

// One leaf step: test the triangle(s) of leaf reference `node`, then pop (or stop: a settled shadow ray).
template <bool COUNT>
__device__ __forceinline__ void leaf_step(const TriGeom* __restrict__ tris, int& node, const RayCtx& r, Hit& best, int* stack,
                                          int& sp, unsigned& n_tris, const Occluder oc) {
    const bool settled = leaf_test<COUNT>(tris, node, r, best, n_tris, oc);
    node = settled ? MCPT_SENT : stack[--sp];
}

// Closest hit of one ray, one ray per thread from start to finish (megakernel and the batch kernel of seam 2).
template <bool COUNT>
__device__ __forceinline__ Hit trace_closest(const DeviceScene& sc, const float3 o, const float3 d, const float tmax,
                                             unsigned& n_nodes, unsigned& n_tris, const Occluder oc = Occluder{-1.0f, -1}) {
    Hit best; best.t = tmax; best.u = 0.f; best.v = 0.f; best.k = -1;
    if (d.x == 0.0f) return best;
    const RayCtx r = ray_setup(o, d, sc.qgrid);
    int stack[MCPT_STACK_SIZE + 4];
    stack[0] = MCPT_SENT;
    int sp = 1;
    int node = sc.root;
    while (node != MCPT_SENT) {
#if MCPT_WIDE
        if (node >= 0) node4_step<COUNT>(sc.qnodes4, node, r, best.t, LocalStack{stack}, sp, n_nodes);
#else
        if (node >= 0) node_step<COUNT>(sc.qnodes, node, r, best.t, stack, sp, n_nodes);
#endif
        else leaf_step<COUNT>(sc.tri, node, r, best, stack, sp, n_tris, oc);
    }
    return best;
}

}  // namespace mcpt
