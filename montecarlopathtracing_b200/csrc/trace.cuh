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

#define MCPT_STACK_SIZE 64
// Edge tolerance of the FP32 barycentric test.  The reference's FP64 same-side test is inclusive (>= 0), so a ray
// through a shared edge hits one of the two triangles; in FP32 the rounding of u, v can put it just outside BOTH
// (a crack: the image diagonals of cornell-box run exactly along wall edges).  Accepting |excess| <= 2e-6 closes the
// cracks; which of two edge-sharing triangles wins is then decided by t like in the reference.
#define MCPT_EDGE_EPS 2.0e-6f

template <bool COUNT>
__device__ __forceinline__ Hit trace_closest(const DeviceScene& sc, const float3 o, const float3 d, const float tmax,
                                             unsigned& n_nodes, unsigned& n_tris) {
    Hit best; best.t = tmax; best.u = 0.f; best.v = 0.f; best.k = -1;
    if (d.x == 0.0f) return best;
    const float ooeps = 1.0e-24f;  // avoid inf/NaN slabs for axis-parallel rays
    const float idx = 1.0f / (fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x));
    const float idy = 1.0f / (fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y));
    const float idz = 1.0f / (fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z));
    const float oodx = o.x * idx, oody = o.y * idy, oodz = o.z * idz;
    const BvhNode* __restrict__ nodes = sc.nodes;
    const TriGeom* __restrict__ tris = sc.tri;
    int stack[MCPT_STACK_SIZE];
    int sp = 0;
    int node = sc.root;
    while (true) {
        if (node >= 0) {
            const float4 n0 = __ldg(&nodes[node].q[0]);
            const float4 n1 = __ldg(&nodes[node].q[1]);
            const float4 nz = __ldg(&nodes[node].q[2]);
            const float4 cq = __ldg(&nodes[node].q[3]);
            if (COUNT) n_nodes++;
            const float a0 = fmaf(n0.x, idx, -oodx), b0 = fmaf(n0.y, idx, -oodx);
            const float a1 = fmaf(n0.z, idy, -oody), b1 = fmaf(n0.w, idy, -oody);
            const float a2 = fmaf(nz.x, idz, -oodz), b2 = fmaf(nz.y, idz, -oodz);
            const float c0min = fmaxf(fmaxf(fminf(a0, b0), fminf(a1, b1)), fmaxf(fminf(a2, b2), 0.0f));
            const float c0max = fminf(fminf(fmaxf(a0, b0), fmaxf(a1, b1)), fminf(fmaxf(a2, b2), best.t));
            const float e0 = fmaf(n1.x, idx, -oodx), f0 = fmaf(n1.y, idx, -oodx);
            const float e1 = fmaf(n1.z, idy, -oody), f1 = fmaf(n1.w, idy, -oody);
            const float e2 = fmaf(nz.z, idz, -oodz), f2 = fmaf(nz.w, idz, -oodz);
            const float c1min = fmaxf(fmaxf(fminf(e0, f0), fminf(e1, f1)), fmaxf(fminf(e2, f2), 0.0f));
            const float c1max = fminf(fminf(fmaxf(e0, f0), fmaxf(e1, f1)), fminf(fmaxf(e2, f2), best.t));
            const bool h0 = c0min <= c0max, h1 = c1min <= c1max;
            const int c0 = __float_as_int(cq.x), c1 = __float_as_int(cq.y);
            if (h0 && h1) {
                const bool swap = c1min < c0min;
                node = swap ? c1 : c0;
                if (sp < MCPT_STACK_SIZE) stack[sp++] = swap ? c0 : c1;
                continue;
            }
            if (h0) { node = c0; continue; }
            if (h1) { node = c1; continue; }
        } else {
            const int k = ~node;
            const float4 v0 = __ldg(&tris[k].v0m);
            const float4 e1 = __ldg(&tris[k].e1f);
            const float4 e2 = __ldg(&tris[k].e2);
            if (COUNT) n_tris++;
            const float3 E1 = f3(e1), E2 = f3(e2);
            const float3 pvec = cross(d, E2);
            const float inv_det = 1.0f / dot(E1, pvec);  // det == 0 -> inf/NaN -> every comparison below fails
            const float3 tvec = o - f3(v0);
            const float u = dot(tvec, pvec) * inv_det;
            const float3 qvec = cross(tvec, E1);
            const float v = dot(d, qvec) * inv_det;
            const float t = dot(E2, qvec) * inv_det;
            if (u >= -MCPT_EDGE_EPS && v >= -MCPT_EDGE_EPS && u + v <= 1.0f + MCPT_EDGE_EPS && t > 0.0f && (t < best.t || (t == best.t && k < best.k))) {
                // quirk Q2 in full: the reference keeps a hit only if t = (p.x - o.x)/d.x > 0 with p.x = o.x + d.x*t formed
                // in FP64 (MTPC/pathTracing.cpp:347,355); when |d.x*t| is below half an ulp of o.x that is 0 and the
                // hit is dropped.  Candidates that pass the FP32 test are rare (1-2 per ray), so the FP64 check is cheap.
                if (__dadd_rn((double)o.x, __dmul_rn((double)d.x, (double)t)) != (double)o.x) {
                    best.t = t; best.u = u; best.v = v; best.k = k;
                }
            }
        }
        if (sp == 0) break;
        node = stack[--sp];
    }
    return best;
}

}  // namespace mcpt
