// render.cu — seams 2 and 3 on the GPU: the batch closest-hit kernel and the path-tracing
// integrator (megakernel with in-register path regeneration).  The integrator is the iterative form
// of the reference's recursive shade() (MTPC/pathTracing.cpp:137-266) with every quirk of SURVEY.md
// 3A kept: un-normalised shading normals (Q3), one un-jittered primary ray per pixel (Q4),
// u->row / v->column texture lookup (Q5), light-selection range frozen to light 0's area (Q6),
// centroid-biased light points (Q7), the double receiver cosine and max(1,d) clamp (Q8), lobe
// pick without pdf division (Q9).
#include "render.h"
#include "philox.cuh"
#include "trace.cuh"

namespace mcpt {

#define MCPT_PI 3.1415926f       /* MTPC/pathTracing.h:11 */
#define MCPT_P_RR 0.6f           /* MTPC/pathTracing.cpp:237 */
#define MCPT_MAX_DEPTH 64        /* reference recursion is unbounded; 0.6^64 ~ 6e-15 (same cap as the oracle) */
enum { LOBE_DIFFUSE = 0, LOBE_SPECULAR = 1, LOBE_TRANSMISSION = 2 };

struct PathVertex { float3 p, pn, kd, wo; int mat; };

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

// One light's next-event estimate (MTPC/pathTracing.cpp:171-231).  Returns false when nothing can be
// added (no triangle selected, receiver faces away, zero weight) so the shadow ray is not traced.
// contrib = kd/pi * Le * |w.vn|/|vn| * |w.pn|/|pn| * A / max(1,d)^2 * (w.pn); the caller multiplies by V.
__device__ __forceinline__ bool sample_light(const DeviceScene& sc, const DLight& L, const PathVertex& vx, const float4 u,
                                             float3& ray_o, float3& ray_d, float& ray_tmax, float3& contrib) {
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
    ray_tmax = (dist - 0.01f) * 1.0002f;  // the sampled triangle itself bounds the closest hit
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
            ray_o = vx.p;  // no origin offset on this branch (:102,:109)
            weight = f3(1.0f, 1.0f, 1.0f);
            if (refract_dir(inc, normal, n1 / n2, ray_d)) { lobe = LOBE_TRANSMISSION; return; }
            ray_d = inc - normal * (2.0f * dot(inc, normal));
            lobe = LOBE_SPECULAR; weight = ks;
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

// ---- K6: megakernel ------------------------------------------------------------------------------------
// One thread per (pixel, group of sample chunks); a warp covers an 8x4 pixel tile.  The thread walks
// its samples in order and every loop iteration traces exactly ONE ray through the single traversal
// call site — the primary ray, a shadow ray of light `phase`, or the extend ray — so lanes whose path
// ended regenerate immediately and the warp stays converged on the traversal loop.  Per-chunk sums
// are accumulated in sample order in a register and written once: no float atomics, so the image is
// a pure function of (seed, pixel, sample) and of the fixed chunk tree (DESIGN.md "determinism").
enum { PH_PRIMARY = -2, PH_NEWSAMPLE = -1 };

template <bool COUNT, bool REUSE_PRIMARY>
__global__ void __launch_bounds__(128) pt_megakernel(const RenderParams P) {
    const DeviceScene& sc = P.scene;
    const int W = P.cam.width, H = P.cam.height;
    const int tiles_x = (W + 7) >> 3, tiles_y = (H + 3) >> 2, n_tiles = tiles_x * tiles_y;
    const int n_groups = P.n_chunks / P.chunks_per_thread;
    const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp_id >= (long long)n_tiles * n_groups) return;
    const int lane = threadIdx.x & 31;
    const int group = (int)(warp_id / n_tiles), tile = (int)(warp_id % n_tiles);
    const int x = (tile % tiles_x) * 8 + (lane & 7), y = (tile / tiles_x) * 4 + (lane >> 3);
    unsigned n_rays = 0, n_nodes = 0, n_tris = 0;
    if (x < W && y < H) {
        const unsigned pixel = (unsigned)(y * W + x);
        const size_t plane_stride = (size_t)W * H * 3;
        // camera ray through the pixel's top-left corner (MTPC/pathTracing.cpp:294-308), FP64 then narrowed
        float3 cam_o, cam_d;
        {
            const double px = P.cam.start[0] - P.cam.pdy[0] * y + P.cam.pdx[0] * x - P.cam.eye[0];
            const double py = P.cam.start[1] - P.cam.pdy[1] * y + P.cam.pdx[1] * x - P.cam.eye[1];
            const double pz = P.cam.start[2] - P.cam.pdy[2] * y + P.cam.pdx[2] * x - P.cam.eye[2];
            const double r = 1.0 / sqrt(px * px + py * py + pz * pz);
            cam_d = f3((float)(px * r), (float)(py * r), (float)(pz * r));
            cam_o = f3((float)P.cam.eye[0], (float)P.cam.eye[1], (float)P.cam.eye[2]);
        }
        const int nl = sc.n_lights;
        int chunk = P.chunk_begin + group * P.chunks_per_thread;       // global chunk index
        const int chunk_end = chunk + P.chunks_per_thread;
        int s = (int)((long long)chunk * P.spp / MCPT_NUM_CHUNKS);
        int s_end = (int)((long long)(chunk + 1) * P.spp / MCPT_NUM_CHUNKS);
        float3 sum = f3(0.f, 0.f, 0.f);
        Hit hit0; hit0.k = -1; hit0.t = 0.f; hit0.u = 0.f; hit0.v = 0.f;
        bool have_hit0 = false;
        PathVertex vx; vx.mat = 0; vx.p = vx.pn = vx.kd = vx.wo = f3(0.f, 0.f, 0.f);
        float3 beta = f3(1.f, 1.f, 1.f), L = f3(0.f, 0.f, 0.f), pend = f3(0.f, 0.f, 0.f), wnext = f3(1.f, 1.f, 1.f);
        float3 ro = cam_o, rd = cam_d; float rtmax = 3.0e38f;
        int phase = PH_NEWSAMPLE, lobe = LOBE_DIFFUSE, light_mat = -1;
        unsigned bounce = 0;
        bool done = false;

        // sample finished with radiance Ls: add in sample order, close chunks as they fill
        auto finish_sample = [&](const float3 Ls) { sum = sum + Ls; s++; phase = PH_NEWSAMPLE; };
        // start shading at a (primary or extend) hit that is not a light
        auto begin_path = [&](const Hit& h) {
            if (h.k < 0) { finish_sample(f3(0.f, 0.f, 0.f)); return; }
            const int m = __float_as_int(__ldg(&sc.tri[h.k].v0m).w);
            const int li = sc.materials[m].light_id;
            if (li >= 0) { finish_sample(f3(sc.lights[li].radiance[0], sc.lights[li].radiance[1], sc.lights[li].radiance[2])); return; }
            make_vertex(sc, h, cam_d, vx);
            beta = f3(1.f, 1.f, 1.f); L = f3(0.f, 0.f, 0.f); bounce = 0; phase = 0;
        };

        while (true) {
            // ---- generate the next ray of this lane (may run through several ray-less states) ----
            bool have_ray = false;
            while (!have_ray) {
                if (phase == PH_NEWSAMPLE) {
                    while (s >= s_end && chunk < chunk_end) {   // chunk complete (or empty): write its plane
                        float* out = P.planes + (size_t)(chunk - P.chunk_begin) * plane_stride + (size_t)pixel * 3;
                        out[0] = sum.x; out[1] = sum.y; out[2] = sum.z;
                        sum = f3(0.f, 0.f, 0.f);
                        chunk++;
                        s_end = (int)((long long)(chunk + 1) * P.spp / MCPT_NUM_CHUNKS);
                    }
                    if (chunk >= chunk_end) { done = true; break; }
                    if (REUSE_PRIMARY && have_hit0) begin_path(hit0);   // identical primary ray (quirk Q4): reuse its hit
                    else { phase = PH_PRIMARY; ro = cam_o; rd = cam_d; rtmax = 3.0e38f; have_ray = true; }
                } else if (phase < nl) {
                    const float4 u = philox_uniform4(bounce, STREAM_LIGHT0 + phase, P.seed_lo, P.seed_hi, pixel, (unsigned)s);
                    const DLight& Lt = sc.lights[phase];
                    if (sample_light(sc, Lt, vx, u, ro, rd, rtmax, pend)) { light_mat = Lt.mat_id; have_ray = true; }
                    else phase++;
                } else {
                    const float4 u = philox_uniform4(bounce, STREAM_PATH, P.seed_lo, P.seed_hi, pixel, (unsigned)s);
                    if (u.x < MCPT_P_RR && bounce + 1 < MCPT_MAX_DEPTH) {   // russian_Roulette (:3-11,:240)
                        next_ray(sc, vx, u, bounce, P.seed_lo, P.seed_hi, pixel, (unsigned)s, ro, rd, lobe, wnext);
                        rtmax = 3.0e38f;
                        have_ray = true;
                    } else finish_sample(L);
                }
            }
            if (done) break;
            // ---- the single traversal site ----
            n_rays++;
            const Hit h = trace_closest<COUNT>(sc, ro, rd, rtmax, n_nodes, n_tris);
            // ---- consume the result ----
            if (phase == PH_PRIMARY) {
                if (REUSE_PRIMARY) { hit0 = h; have_hit0 = true; }
                begin_path(h);
            } else if (phase < nl) {
                // visible iff the CLOSEST hit carries the light's material (:212-215)
                if (h.k >= 0 && __float_as_int(__ldg(&sc.tri[h.k].v0m).w) == light_mat) L = L + beta * pend;
                phase++;
            } else {
                if (h.k < 0) { finish_sample(L); continue; }
                const int m = __float_as_int(__ldg(&sc.tri[h.k].v0m).w);
                const int li = sc.materials[m].light_id;
                const float3 wgt = wnext * (1.0f / MCPT_P_RR);
                if (li >= 0) {   // diffuse rays drop emitters (NEE counted them, :247-253); specular/transmission keep them
                    if (lobe != LOBE_DIFFUSE)
                        L = L + beta * wgt * f3(sc.lights[li].radiance[0], sc.lights[li].radiance[1], sc.lights[li].radiance[2]);
                    finish_sample(L);
                } else {
                    beta = beta * wgt;
                    make_vertex(sc, h, rd, vx);
                    bounce++;
                    phase = 0;
                }
            }
        }
    }
    // device-side ray accounting: one atomic per warp
    const unsigned mask = __activemask();
    const unsigned r = __reduce_add_sync(mask, n_rays);
    if (lane == __ffs(mask) - 1) {
        atomicAdd(&P.counters[0], (unsigned long long)r);
    }
    if (COUNT) {
        const unsigned a = __reduce_add_sync(mask, n_nodes), b = __reduce_add_sync(mask, n_tris);
        if (lane == __ffs(mask) - 1) { atomicAdd(&P.counters[1], (unsigned long long)a); atomicAdd(&P.counters[2], (unsigned long long)b); }
    }
}

// ---- K5: batch closest hit (parity hook of seam 2 and the traversal-only benchmark) -----------------
template <bool COUNT>
__global__ void __launch_bounds__(128) trace_kernel(const DeviceScene sc, const float* __restrict__ org_dir, const int n,
                                                    mcpt_hit* __restrict__ hits, unsigned long long* counters) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned n_nodes = 0, n_tris = 0;
    if (i < n) {
        const float* r = org_dir + 6 * (size_t)i;
        const float3 o = f3(r[0], r[1], r[2]), d = f3(r[3], r[4], r[5]);
        const Hit h = trace_closest<COUNT>(sc, o, d, 3.0e38f, n_nodes, n_tris);
        mcpt_hit out;
        out.t = h.k >= 0 ? h.t : 0.0f; out.u = h.u; out.v = h.v;
        out.leaf = h.k;
        out.tri = h.k >= 0 ? sc.leaf_face[h.k] : -1;
        hits[i] = out;
    }
    if (COUNT) {
        const unsigned mask = __activemask();
        const unsigned a = __reduce_add_sync(mask, n_nodes), b = __reduce_add_sync(mask, n_tris);
        if ((threadIdx.x & 31) == __ffs(mask) - 1) { atomicAdd(&counters[1], (unsigned long long)a); atomicAdd(&counters[2], (unsigned long long)b); }
    }
}

// ---- K12: fixed pairwise tree over sample-chunk planes ------------------------------------------------
// ((p0+p1)+(p2+p3))+((p4+p5)+(p6+p7)): the same association whether the planes were produced by one GPU
// or gathered from 2/4/8, so the image does not depend on the GPU count.
template <int N>
__global__ void __launch_bounds__(256) sum_planes_kernel(const float* __restrict__ planes, const long long count, const float scale,
                                                         float* __restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        float v[N];
#pragma unroll
        for (int k = 0; k < N; k++) v[k] = __ldcs(planes + (size_t)k * count + i);
#pragma unroll
        for (int w = 1; w < N; w *= 2)
#pragma unroll
            for (int k = 0; k < N; k += 2 * w) v[k] = __fadd_rn(v[k], v[k + w]);
        out[i] = __fmul_rn(v[0], scale);
    }
}

int launch_trace(const DeviceScene& sc, const float* d_org_dir, int n, mcpt_hit* d_hits, bool count,
                 unsigned long long* d_counters, cudaStream_t stream) {
    if (n <= 0) return 0;
    const int tpb = 128, blocks = (n + tpb - 1) / tpb;
    if (count) trace_kernel<true><<<blocks, tpb, 0, stream>>>(sc, d_org_dir, n, d_hits, d_counters);
    else trace_kernel<false><<<blocks, tpb, 0, stream>>>(sc, d_org_dir, n, d_hits, d_counters);
    return (int)cudaGetLastError();
}

int launch_megakernel(const RenderParams& P, int flags, cudaStream_t stream) {
    const int W = P.cam.width, H = P.cam.height;
    const long long n_tiles = (long long)((W + 7) / 8) * ((H + 3) / 4);
    const long long warps = n_tiles * (P.n_chunks / P.chunks_per_thread);
    const int tpb = 128;
    const long long blocks = (warps * 32 + tpb - 1) / tpb;
    if (blocks <= 0 || blocks > 0x7fffffffLL) return (int)cudaErrorInvalidConfiguration;
    const bool count = flags & MCPT_FLAG_COUNT, reuse = !(flags & MCPT_FLAG_NO_PRIMARY_REUSE);
    if (count && reuse) pt_megakernel<true, true><<<(unsigned)blocks, tpb, 0, stream>>>(P);
    else if (count) pt_megakernel<true, false><<<(unsigned)blocks, tpb, 0, stream>>>(P);
    else if (reuse) pt_megakernel<false, true><<<(unsigned)blocks, tpb, 0, stream>>>(P);
    else pt_megakernel<false, false><<<(unsigned)blocks, tpb, 0, stream>>>(P);
    return (int)cudaGetLastError();
}

int launch_sum_planes(const float* d_planes, int n, long long count, float scale, float* d_out, cudaStream_t stream) {
    if (count <= 0) return 0;
    const int tpb = 256;
    long long want = (count + tpb - 1) / tpb;
    const int blocks = (int)(want < 148 * 8 ? want : 148 * 8);  // grid-stride, 8 CTAs per SM
    switch (n) {
        case 1: sum_planes_kernel<1><<<blocks, tpb, 0, stream>>>(d_planes, count, scale, d_out); break;
        case 2: sum_planes_kernel<2><<<blocks, tpb, 0, stream>>>(d_planes, count, scale, d_out); break;
        case 4: sum_planes_kernel<4><<<blocks, tpb, 0, stream>>>(d_planes, count, scale, d_out); break;
        case 8: sum_planes_kernel<8><<<blocks, tpb, 0, stream>>>(d_planes, count, scale, d_out); break;
        default: return (int)cudaErrorInvalidValue;
    }
    return (int)cudaGetLastError();
}

}  // namespace mcpt
