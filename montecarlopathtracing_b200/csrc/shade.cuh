// shade.cuh — the per-vertex arithmetic of the integrator, shared by the megakernel (render.cu) and the wavefront
// kernels (wavefront.cu) so that both produce the same floating-point results for the same Philox keys.
// Iterative form of the reference's recursive shade() (MTPC/pathTracing.cpp:137-266) with every quirk of SURVEY.md 3A
// kept: un-normalised shading normals (Q3), u->row / v->column texture lookup (Q5), light-selection range frozen to
// light 0's area (Q6), centroid-biased light points (Q7), the double receiver cosine and max(1,d) clamp (Q8), lobe
// pick without pdf division (Q9).
#pragma once
#include "device_types.cuh"
#include "exact.cuh"
#include "philox.cuh"
#include "trace.cuh"
#include "vecmath.cuh"

namespace mcpt {

// explicitly rounded products/sums for the radiance bookkeeping: no FMA contraction, so the megakernel and the
// wavefront kernels accumulate bit-identically whatever the surrounding code looks like
__device__ __forceinline__ float3 mul_rn(float3 a, float3 b) { return f3(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y), __fmul_rn(a.z, b.z)); }
__device__ __forceinline__ float3 add_rn(float3 a, float3 b) { return f3(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z)); }

#define MCPT_PI 3.1415926f       /* MTPC/pathTracing.h:11 */
#define MCPT_P_RR 0.6f           /* MTPC/pathTracing.cpp:237 */
#define MCPT_SELF_EPS 1.0e-4f    /* 1 % of the reference's 0.01 ray offset */
#define MCPT_MAX_DEPTH 64        /* reference recursion is unbounded; 0.6^64 ~ 6e-15 (same cap as the oracle) */
enum { LOBE_DIFFUSE = 0, LOBE_SPECULAR = 1, LOBE_TRANSMISSION = 2 };

struct PathVertex { float3 p, pn, kd, wo; int mat; int k; float t; };   // k = sorted-leaf index of the vertex's triangle, t = length of the ray that found it

// Hit -> shading point: position and barycentric normal (MTPC/pathTracing.cpp:345-351, not normalised),
// material, kd or texel (MTPC/pathTracing.cpp:148-163).
__device__ __forceinline__ void make_vertex(const DeviceScene& sc, const Hit& h, const float3 ray_d, PathVertex& vx) {
    const TriGeom* g = &sc.tri[h.k];
    const float4 v0 = __ldg(&g->v0m), e1 = __ldg(&g->e1f), e2 = __ldg(&g->e2);
    const TriShade* s = &sc.shade[h.k];
    const float4 sa = __ldg(&s->a), sb = __ldg(&s->b), sc4 = __ldg(&s->c), sd = __ldg(&s->d);
    const float w = 1.0f - h.u - h.v;
    vx.p = f3(v0) + f3(e1) * h.u + f3(e2) * h.v;
    vx.pn = f3(sa.x, sa.y, sa.z) * w + f3(sa.w, sb.x, sb.y) * h.u + f3(sb.z, sb.w, sc4.x) * h.v;
    vx.wo = -ray_d;
    vx.mat = __float_as_int(v0.w);
    vx.k = h.k; vx.t = h.t;
    const DMaterial* m = &sc.materials[vx.mat];
    if (m->tex_id >= 0) {
        const DTexture tx = sc.textures[m->tex_id];
        const float row = sc4.y * w + sc4.w * h.u + sd.y * h.v;   // vt.x -> row (quirk Q5)
        const float col = sc4.z * w + sd.x * h.u + sd.z * h.v;    // vt.y -> column
        const float irow = row - floorf(row), icol = col - floorf(col);
        const int r = min((int)(irow * tx.height), tx.height - 1), c = min((int)(icol * tx.width), tx.width - 1);
        const unsigned char* px = sc.tex_bytes + tx.offset + ((size_t)r * tx.width + c) * 3;
        vx.kd = f3(px[2] * (1.0f / 255.0f), px[1] * (1.0f / 255.0f), px[0] * (1.0f / 255.0f));  // BGR -> RGB
    } else {
        vx.kd = f3(m->kd[0], m->kd[1], m->kd[2]);
    }
}

// MCPT_FLAG_JITTER: the primary ray of sample s of pixel (x, y) goes through (x + u, y + v) of the pixel grid, (u, v) two Philox
// uniforms of their own stream; formed in FP64 like raygen_kernel's corner rays (generateImg, MTPC/pathTracing.cpp:276-308).
__device__ __forceinline__ float3 jittered_primary_dir(const DCamera& cam, const int x, const int y, const unsigned seed_lo,
                                                       const unsigned seed_hi, const unsigned pixel, const unsigned sample) {
    const float4 u = philox_uniform4(0u, STREAM_JITTER, seed_lo, seed_hi, pixel, sample);
    const double fx = (double)x + (double)u.x, fy = (double)y + (double)u.y;
    const double dx = cam.start[0] - cam.pdy[0] * fy + cam.pdx[0] * fx - cam.eye[0];
    const double dy = cam.start[1] - cam.pdy[1] * fy + cam.pdx[1] * fx - cam.eye[1];
    const double dz = cam.start[2] - cam.pdy[2] * fy + cam.pdx[2] * fx - cam.eye[2];
    const double l = sqrt(dx * dx + dy * dy + dz * dz);
    return f3((float)(dx / l), (float)(dy / l), (float)(dz / l));
}

// One light's next-event estimate (MTPC/pathTracing.cpp:171-231).  Returns false when nothing can be
// added (no triangle selected, receiver faces away, zero weight) so the shadow ray is not traced.
// contrib = kd/pi * Le * |w.vn|/|vn| * |w.pn|/|pn| * A / max(1,d)^2 * (w.pn); the caller multiplies by V.
__device__ __forceinline__ bool sample_light(const DeviceScene& sc, const DLight& L, const PathVertex& vx, const float4 u,
                                             float3& ray_o, float3& ray_d, float& ray_tmax, float3& contrib, const bool anyhit = false) {
    const float rnd = u.x * sc.area0;  // quirk Q6
    const float* cdf = sc.light_cdf + L.first;
    int lo = 0, hi = L.n_search;      // first j with rnd < cdf[j]
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (rnd < __ldg(&cdf[mid])) hi = mid; else lo = mid + 1; }
    if (lo >= L.count) return false;  // nothing selected (reference: NaN); documented deviation
    const LightTri* lt = &sc.light_tris[L.first + lo];
    const float4 A = __ldg(&lt->p0n0x), B = __ldg(&lt->p1n0y), C = __ldg(&lt->p2n0z), D = __ldg(&lt->n1n2x), E = __ldg(&lt->n2yz);
    const float inv = 1.0f / (u.y + u.z + u.w);
    const float p1 = u.y * inv, p2 = u.z * inv, p3 = u.w * inv;  // quirk Q7
    const float3 xl = f3(A) * p1 + f3(B) * p2 + f3(C) * p3;
    const float3 vn = f3(A.w, B.w, C.w) * p1 + f3(D.x, D.y, D.z) * p2 + f3(D.w, E.x, E.y) * p3;
    const float3 dv = xl - vx.p;
    const float dist = length(dv);
    const float3 w = dv * (1.0f / dist);
    const float cosp = dot(w, vx.pn);
    if (!(cosp > 0.0f)) return false;
    const float cos_theta = fabsf(dot(w, vn)) / length(vn);
    const float cos_hat = fabsf(cosp) / length(vx.pn);
    const float dc = fmaxf(1.0f, dist);
    const float g = cos_theta * cos_hat / (dc * dc) * L.area * cosp * (1.0f / MCPT_PI);  // quirk Q8
    contrib = vx.kd * f3(L.radiance[0], L.radiance[1], L.radiance[2]) * g;
    if (!(contrib.x != 0.0f || contrib.y != 0.0f || contrib.z != 0.0f)) return false;
    ray_o = vx.p + w * 0.01f;
    ray_d = w;
    // closest-hit rule: the sampled triangle itself bounds the closest hit; any-hit rule: stop short of the sampled point
    ray_tmax = (dist - 0.01f) * (anyhit ? 0.9999f : 1.0002f);
    return true;
}

// BRDFImportanceSampling (MTPC/pathTracing.cpp:30-64) around the un-normalised axis.  sin/cos of theta
// are formed without the asin/acos round trip: diffuse theta = asin(sqrt(u)); specular
// cos(theta) = u^(1/(Ns+1)) evaluated as 1 + expm1(log(u)/(Ns+1)) so that Ns = 1e5 keeps its precision.
__device__ __forceinline__ float3 brdf_sample(const float3 axis, const int lobe, const float Ns, const float u_phi, const float u_theta) {
    float sphi, cphi;
    sincosf(u_phi * 2.0f * MCPT_PI, &sphi, &cphi);
    float st, ct;
    if (lobe == LOBE_DIFFUSE) { st = sqrtf(u_theta); ct = sqrtf(1.0f - u_theta); }
    else { const float om = -expm1f(logf(u_theta) / (Ns + 1.0f)); ct = 1.0f - om; st = sqrtf(om * (2.0f - om)); }
    float3 front;
    if (fabsf(axis.x) > fabsf(axis.y)) front = normalize(f3(axis.z, 0.0f, -axis.x));
    else front = normalize(f3(0.0f, -axis.z, axis.y));
    const float3 right = cross(axis, front);
    return normalize(right * (st * cphi) + axis * ct + front * (st * sphi));
}

// Refract (MTPC/pathTracing.cpp:13-27)
__device__ __forceinline__ bool refract_dir(const float3 i, const float3 n, const float eta, float3& out) {
    const float cosi = dot(i, n);
    const float cost2 = 1.0f - eta * eta * (1.0f - cosi * cosi);
    if (cost2 >= 0.0f) { out = i * eta - n * (eta * cosi + sqrtf(cost2)); return true; }
    return false;
}

// nextRay (MTPC/pathTracing.cpp:66-134).  u = (rr, lobe, phi, theta) of STREAM_PATH.
__device__ __forceinline__ void next_ray(const DeviceScene& sc, const PathVertex& vx, const float4 u, const unsigned bounce,
                                         const unsigned seed_lo, const unsigned seed_hi, const unsigned pixel, const unsigned sample,
                                         float3& ray_o, float3& ray_d, int& lobe, float3& weight) {
    const DMaterial* m = &sc.materials[vx.mat];
    const float3 ks = f3(m->ks[0], m->ks[1], m->ks[2]);
    const float3 inc = -vx.wo;
    if (m->Ni > 1.0f) {
        const float cos_in = dot(inc, vx.pn);
        float3 normal; float n1, n2;
        if (cos_in > 0.0f) { normal = -vx.pn; n1 = m->Ni; n2 = 1.0f; } else { normal = vx.pn; n1 = 1.0f; n2 = m->Ni; }
        const float r = (n1 - n2) / (n1 + n2), rf0 = r * r;
        const float a = 1.0f - fabsf(cos_in);
        const float fresnel = rf0 + (1.0f - rf0) * (a * a * a * a * a);
        const float4 uf = philox_uniform4(bounce, STREAM_FRESNEL, seed_lo, seed_hi, pixel, sample);
        if (fresnel < uf.x) {
            weight = f3(1.0f, 1.0f, 1.0f);
            if (refract_dir(inc, normal, n1 / n2, ray_d)) lobe = LOBE_TRANSMISSION;
            else { ray_d = inc - normal * (2.0f * dot(inc, normal)); lobe = LOBE_SPECULAR; weight = ks; }
            // no origin offset on this branch (:102,:109): whether the reference's ray hits its own triangle again is a property
            // of its FP64 arithmetic (exact.cuh).  Reproduce the verdict, then start a hair behind the surface (the triangle is
            // hit again, at the same point) or a hair in front of it (it is not), which FP32 traversal resolves reliably.
            const bool again = reference_self_hit_fp64(sc.pos, sc.leaf_face[vx.k], vx.p, inc, vx.t, ray_d);
            ray_o = vx.p + ray_d * (again ? -MCPT_SELF_EPS : MCPT_SELF_EPS);
            return;
        }
    }
    const float kd_norm = length(vx.kd);
    if (m->ks_norm != 0.0f && kd_norm / m->ks_norm < u.y) {  // quirk Q9
        const float3 reflect = inc - vx.pn * (2.0f * dot(inc, vx.pn));
        ray_d = brdf_sample(reflect, LOBE_SPECULAR, m->Ns, u.z, u.w);
        lobe = LOBE_SPECULAR; weight = ks;
    } else {
        ray_d = brdf_sample(vx.pn, LOBE_DIFFUSE, m->Ns, u.z, u.w);
        lobe = LOBE_DIFFUSE; weight = vx.kd;
    }
    ray_o = vx.p + ray_d * 0.01f;
}

// ---- MCPT_FLAG_MIS_BALANCE: a physically based estimator, NOT the reference's ---------------------------
// The reference partitions light transport by lobe — next-event estimation for the diffuse lobe only, BSDF sampling for the
// specular lobe only, no pdfs (quirks Q6-Q9) — which is why its veach-mis render is noisy on the plates.  This mode is the
// textbook alternative: a Lambert + normalised-Phong BRDF  f = kd/pi + ks (Ns+2)/(2 pi) cos^Ns(alpha)  on the NORMALISED shading
// normal turned toward the viewer, uniform area sampling of every light (one sample per light and vertex) and BSDF sampling
// (lobe picked with probability |kd| / (|kd| + |ks|)), combined with the balance heuristic; any-hit visibility.  Same Philox
// streams and Russian roulette as the parity mode.  oracle/mcpt_oracle.c has the FP64 twin (tests compare them pixel by pixel).
struct MisFrame { float3 n, r, kd, ks; float pd, Ns; };
__device__ __forceinline__ MisFrame mis_frame(const DeviceScene& sc, const PathVertex& vx) {
    const DMaterial* m = &sc.materials[vx.mat];
    MisFrame F;
    F.n = normalize(vx.pn);
    if (dot(F.n, vx.wo) < 0.0f) F.n = -F.n;
    const float two_c = __fmul_rn(2.0f, dot(F.n, vx.wo));   // explicit roundings: both integrators must form the same bits
    F.r = f3(__fmaf_rn(F.n.x, two_c, -vx.wo.x), __fmaf_rn(F.n.y, two_c, -vx.wo.y), __fmaf_rn(F.n.z, two_c, -vx.wo.z));
    F.kd = vx.kd; F.ks = f3(m->ks[0], m->ks[1], m->ks[2]); F.Ns = m->Ns;
    const float kdn = length(vx.kd), ksn = m->ks_norm;
    F.pd = (kdn + ksn) > 0.0f ? kdn / (kdn + ksn) : 1.0f;
    return F;
}
__device__ __forceinline__ float mis_lobe(const MisFrame& F, const float3 wi) { return powf(fmaxf(dot(F.r, wi), 0.0f), F.Ns); }
__device__ __forceinline__ float3 mis_brdf(const MisFrame& F, const float lobe) {
    const float sp = __fmul_rn(__fmul_rn(__fadd_rn(F.Ns, 2.0f), 0.5f / MCPT_PI), lobe);
    return f3(__fmaf_rn(F.ks.x, sp, __fmul_rn(F.kd.x, 1.0f / MCPT_PI)), __fmaf_rn(F.ks.y, sp, __fmul_rn(F.kd.y, 1.0f / MCPT_PI)),
              __fmaf_rn(F.ks.z, sp, __fmul_rn(F.kd.z, 1.0f / MCPT_PI)));
}
__device__ __forceinline__ float mis_pdf_bsdf(const MisFrame& F, const float cosi, const float lobe) {
    const float sp = __fmul_rn(__fmul_rn(__fmul_rn(__fsub_rn(1.0f, F.pd), __fadd_rn(F.Ns, 1.0f)), 0.5f / MCPT_PI), lobe);
    return __fmaf_rn(__fmul_rn(F.pd, cosi), 1.0f / MCPT_PI, sp);
}
// solid-angle pdf with which light L's area sampling picks the point at distance `dist` along w on a surface of normal ng
__device__ __forceinline__ float mis_pdf_light(const float area, const float dist, const float cosl) {
    return __fdiv_rn(__fmul_rn(dist, dist), __fmul_rn(area, cosl));
}

// contrib = f cos Le / pdf_L * pdf_L / (pdf_L + pdf_B): the caller multiplies by the throughput and the any-hit visibility
__device__ __forceinline__ bool sample_light_mis(const DeviceScene& sc, const DLight& L, const MisFrame& F, const PathVertex& vx, const float4 u,
                                                 float3& ray_o, float3& ray_d, float& ray_tmax, float3& contrib) {
    const float rnd = u.x * L.area;
    const float* cdf = sc.light_cdf + L.first;
    int lo = 0, hi = L.count - 1;     // first j with rnd < cdf[j] (the last triangle catches rnd == area after rounding)
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (rnd < __ldg(&cdf[mid])) hi = mid; else lo = mid + 1; }
    const LightTri* lt = &sc.light_tris[L.first + lo];
    const float3 A = f3(__ldg(&lt->p0n0x)), B = f3(__ldg(&lt->p1n0y)), C = f3(__ldg(&lt->p2n0z));
    const float su = sqrtf(u.y);
    const float b0 = __fsub_rn(1.0f, su), b1 = __fmul_rn(su, __fsub_rn(1.0f, u.z)), b2 = __fmul_rn(su, u.z);
    const float3 xl = f3(__fmaf_rn(C.x, b2, __fmaf_rn(B.x, b1, __fmul_rn(A.x, b0))), __fmaf_rn(C.y, b2, __fmaf_rn(B.y, b1, __fmul_rn(A.y, b0))),
                         __fmaf_rn(C.z, b2, __fmaf_rn(B.z, b1, __fmul_rn(A.z, b0))));   // uniform over the triangle
    const float3 ng = normalize(cross(B - A, C - A));
    const float3 dv = xl - vx.p;
    const float dist = length(dv);
    const float3 w = dv * (1.0f / dist);
    const float cosi = dot(F.n, w), cosl = fabsf(dot(ng, w));
    if (!(cosi > 0.0f) || !(cosl > 1.0e-6f) || !(dist > 0.0f)) return false;
    const float lobe = mis_lobe(F, w);
    const float g = __fdiv_rn(cosi, __fadd_rn(mis_pdf_light(L.area, dist, cosl), mis_pdf_bsdf(F, cosi, lobe)));
    contrib = mis_brdf(F, lobe) * f3(L.radiance[0], L.radiance[1], L.radiance[2]) * g;
    if (!(contrib.x != 0.0f || contrib.y != 0.0f || contrib.z != 0.0f)) return false;
    ray_o = vx.p + w * 0.01f; ray_d = w;
    ray_tmax = (dist - 0.01f) * 0.9999f;
    return true;
}
// BSDF sampling; weight = f cos / pdf_B, pdf_b = pdf_B of the sampled direction (< 0: the reference's delta refraction branch,
// whose emission is counted in full).  Returns false when the sampled direction leaves the hemisphere of n (the path ends).
__device__ __forceinline__ bool next_ray_mis(const DeviceScene& sc, const MisFrame& F, const PathVertex& vx, const float4 u, const unsigned bounce,
                                             const unsigned seed_lo, const unsigned seed_hi, const unsigned pixel, const unsigned sample,
                                             float3& ray_o, float3& ray_d, int& lobe_tag, float3& weight, float& pdf_b) {
    if (sc.materials[vx.mat].Ni > 1.0f) {
        next_ray(sc, vx, u, bounce, seed_lo, seed_hi, pixel, sample, ray_o, ray_d, lobe_tag, weight);
        pdf_b = -1.0f;
        return true;
    }
    if (u.y < F.pd) { ray_d = brdf_sample(F.n, LOBE_DIFFUSE, F.Ns, u.z, u.w); lobe_tag = LOBE_DIFFUSE; }
    else { ray_d = brdf_sample(F.r, LOBE_SPECULAR, F.Ns, u.z, u.w); lobe_tag = LOBE_SPECULAR; }
    const float cosi = dot(F.n, ray_d);
    if (!(cosi > 0.0f)) return false;
    const float lobe = mis_lobe(F, ray_d);
    pdf_b = mis_pdf_bsdf(F, cosi, lobe);
    if (!(pdf_b > 0.0f)) return false;
    weight = mis_brdf(F, lobe) * __fdiv_rn(cosi, pdf_b);
    ray_o = vx.p + ray_d * 0.01f;
    return true;
}
// Balance-heuristic weight of an emitter found by BSDF sampling: hit triangle k (a light of area `area`) at distance t along d.
__device__ __forceinline__ float mis_weight_bsdf_hit(const DeviceScene& sc, const int k, const float3 d, const float t, const float area, const float pdf_b) {
    if (pdf_b < 0.0f) return 1.0f;
    const float3 ng = normalize(cross(f3(__ldg(&sc.tri[k].e1f)), f3(__ldg(&sc.tri[k].e2))));
    const float cosl = fabsf(dot(ng, d));
    if (!(cosl > 1.0e-6f)) return 1.0f;
    return __fdiv_rn(pdf_b, __fadd_rn(pdf_b, mis_pdf_light(area, t, cosl)));
}

}  // namespace mcpt
