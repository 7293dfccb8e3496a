// exact.cuh — the reference's ray/triangle verdict in FP64, operation for operation (no FMA contraction: the reference is built
// for baseline x86-64, every product and sum is rounded).  Used where FP32 cannot decide what the reference decides:
//   * seam 2 (K5, render.cu): which of several near-tied candidates the reference's closest hit is;
//   * the no-offset rays of the refraction branch (shade.cuh next_ray): whether the reference re-hits the triangle it starts on.
#pragma once
#include "device_types.cuh"

namespace mcpt {

struct D3 { double x, y, z; };
__device__ __forceinline__ D3 d3(double x, double y, double z) { D3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ D3 dsub(D3 a, D3 b) { return d3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ double ddot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ D3 dcross(D3 a, D3 b) { return d3(a.y * b.z - b.y * a.z, b.x * a.z - a.x * b.z, a.x * b.y - b.x * a.y); }

__device__ __forceinline__ D3 dsub_rn(D3 a, D3 b) { return d3(__dsub_rn(a.x, b.x), __dsub_rn(a.y, b.y), __dsub_rn(a.z, b.z)); }
__device__ __forceinline__ double ddot_rn(D3 a, D3 b) {   // Vertex::operator*(Vertex): (x*x2 + y*y2) + z*z2
    return __dadd_rn(__dadd_rn(__dmul_rn(a.x, b.x), __dmul_rn(a.y, b.y)), __dmul_rn(a.z, b.z));
}
__device__ __forceinline__ D3 dcross_rn(D3 a, D3 b) {     // Vertex::cross, MTPC/sceneManagement.h:69-75
    return d3(__dsub_rn(__dmul_rn(a.y, b.z), __dmul_rn(b.y, a.z)), __dsub_rn(__dmul_rn(b.x, a.z), __dmul_rn(a.x, b.z)),
              __dsub_rn(__dmul_rn(a.x, b.y), __dmul_rn(b.x, a.y)));
}
// The reference's verdict on (ray, face): true iff the face is hit with t > 0; t_out = its t.
static __device__ __noinline__ bool reference_tri_test_fp64(const double* __restrict__ pos, const int face, const D3 o, const D3 d, double& t_out) {
    const double* q = pos + 9 * (size_t)face;
    const D3 v1 = d3(q[0], q[1], q[2]), v2 = d3(q[3], q[4], q[5]), v3 = d3(q[6], q[7], q[8]);
    D3 n = dcross_rn(dsub_rn(v1, v2), dsub_rn(v3, v1));                       // Face::calNorm
    const double nl = __dsqrt_rn(ddot_rn(n, n));
    n = d3(__ddiv_rn(n.x, nl), __ddiv_rn(n.y, nl), __ddiv_rn(n.z, nl));
    const double t = __ddiv_rn(ddot_rn(dsub_rn(v1, o), n), ddot_rn(n, d));
    const D3 p = d3(__dadd_rn(o.x, __dmul_rn(d.x, t)), __dadd_rn(o.y, __dmul_rn(d.y, t)), __dadd_rn(o.z, __dmul_rn(d.z, t)));
    const D3 ap = dsub_rn(p, v1), bp = dsub_rn(p, v2), cp = dsub_rn(p, v3);
    const D3 ab = dsub_rn(v2, v1), bc = dsub_rn(v3, v2), ca = dsub_rn(v1, v3);
    const double dir1 = ddot_rn(dcross_rn(ab, ap), n), dir2 = ddot_rn(dcross_rn(bc, bp), n), dir3 = ddot_rn(dcross_rn(ca, cp), n);
    const double j1 = __dmul_rn(dir1, dir2), j2 = __dmul_rn(dir1, dir3), j3 = __dmul_rn(dir2, dir3);
    if (!(j1 >= 0.0 && j2 >= 0.0 && j3 >= 0.0)) return false;                 // NaN (parallel ray) -> miss, as in the reference
    t_out = __ddiv_rn(__dsub_rn(p.x, o.x), d.x);                              // MTPC/pathTracing.cpp:347
    return t_out > 0.0;
}


// The refraction branch starts its ray ON the surface, with no origin offset (MTPC/pathTracing.cpp:102,109), so in the reference
// the triangle the ray leaves is a candidate for its own next hit: the plane hit comes out at t ~ +-1e-16, p.x either rounds back
// onto o.x (t = 0: no hit) or not, and t > 0 accepts it — a coin whose bias (~0.37 for unit-sized coordinates; 0.5 if decided in
// FP32) is a property of the FP64 arithmetic.  The image depends on it (a self-hit refracts twice), so it is reproduced rather
// than decided by FP32 noise: re-form the previous hit point p as the reference did — the FP64 plane hit of the incoming ray
// (direction d_in, length t_in: the rounding noise of p grows with the ray's length, and with it the coin's bias: 0.31 for a
// ray of length 1, 0.40 for length 4) — and ask the reference's own test whether the ray (p, d_out) hits the triangle again.
static __device__ __noinline__ bool reference_self_hit_fp64(const double* __restrict__ pos, const int face, const float3 p32, const float3 d_in,
                                                            const float t_in, const float3 d_out) {
    const double* q = pos + 9 * (size_t)face;
    const D3 v1 = d3(q[0], q[1], q[2]), v2 = d3(q[3], q[4], q[5]), v3 = d3(q[6], q[7], q[8]);
    D3 n = dcross_rn(dsub_rn(v1, v2), dsub_rn(v3, v1));
    const double nl = __dsqrt_rn(ddot_rn(n, n));
    n = d3(__ddiv_rn(n.x, nl), __ddiv_rn(n.y, nl), __ddiv_rn(n.z, nl));
    const D3 di = d3(d_in.x, d_in.y, d_in.z);
    const double back = (double)t_in;                                                    // the incoming ray's origin
    const D3 o = d3((double)p32.x - di.x * back, (double)p32.y - di.y * back, (double)p32.z - di.z * back);
    const double t = __ddiv_rn(ddot_rn(dsub_rn(v1, o), n), ddot_rn(n, di));
    const D3 p = d3(__dadd_rn(o.x, __dmul_rn(di.x, t)), __dadd_rn(o.y, __dmul_rn(di.y, t)), __dadd_rn(o.z, __dmul_rn(di.z, t)));
    double t2;
    return reference_tri_test_fp64(pos, face, p, d3(d_out.x, d_out.y, d_out.z), t2);
}

}  // namespace mcpt
