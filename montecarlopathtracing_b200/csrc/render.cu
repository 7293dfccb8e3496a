// render.cu — seams 2 and 3 on the GPU: the batch closest-hit kernel and the path-tracing
// integrator (megakernel with in-register path regeneration).  The integrator is the iterative form
// of the reference's recursive shade() (MTPC/pathTracing.cpp:137-266) with every quirk of SURVEY.md
// 3A kept: un-normalised shading normals (Q3), one un-jittered primary ray per pixel (Q4),
// u->row / v->column texture lookup (Q5), light-selection range frozen to light 0's area (Q6),
// centroid-biased light points (Q7), the double receiver cosine and max(1,d) clamp (Q8), lobe
// pick without pdf division (Q9).
#include "render.h"
#include "exact.cuh"
#include "shade.cuh"

namespace mcpt {

// ---- K7: primary rays --------------------------------------------------------------------------------
// generateImg walks a row by REPEATED ADDITION: pos = start - pdy*i, then pos += pdx per pixel
// (MTPC/pathTracing.cpp:297,326).  The rounding of that walk is visible: on the centre column of an even-width image
// start.x + j*pdx.x is exactly 0 when formed by one multiplication, which by quirk Q2 (d.x == 0 hits nothing) would
// paint a black column the reference does not have.  The walk is a serial chain per row, so it runs one thread per row —
// but only the walk (K7a: three explicitly rounded FP64 additions per pixel, positions to a scratch array); normalize(pos - eye),
// two dozen times the work (FP64 sqrt and three divisions), runs one thread per PIXEL (K7b).  Same operations in the same order
// as the single kernel of round 1, which spent 0.45 ms on 1024 x 1024 (1.7 ms at 3840 x 2160) with 1 024 threads on the GPU.
__global__ void __launch_bounds__(32) raygen_walk_kernel(const DCamera cam, double* __restrict__ pos) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cam.height) return;
    double px = __dsub_rn(cam.start[0], __dmul_rn(cam.pdy[0], (double)i));
    double py = __dsub_rn(cam.start[1], __dmul_rn(cam.pdy[1], (double)i));
    double pz = __dsub_rn(cam.start[2], __dmul_rn(cam.pdy[2], (double)i));
    double* out = pos + 3 * (size_t)i * cam.width;
    for (int j = 0; j < cam.width; j++) {
        out[3 * j + 0] = px; out[3 * j + 1] = py; out[3 * j + 2] = pz;
        px = __dadd_rn(px, cam.pdx[0]); py = __dadd_rn(py, cam.pdx[1]); pz = __dadd_rn(pz, cam.pdx[2]);
    }
}
__global__ void __launch_bounds__(256) raygen_dir_kernel(const DCamera cam, const double* __restrict__ pos, float* __restrict__ dirs) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (size_t)cam.width * cam.height) return;
    const double dx = __dsub_rn(pos[3 * t + 0], cam.eye[0]), dy = __dsub_rn(pos[3 * t + 1], cam.eye[1]), dz = __dsub_rn(pos[3 * t + 2], cam.eye[2]);
    const double l = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)));
    dirs[3 * t + 0] = (float)__ddiv_rn(dx, l);
    dirs[3 * t + 1] = (float)__ddiv_rn(dy, l);
    dirs[3 * t + 2] = (float)__ddiv_rn(dz, l);
}

// d_pos: scratch of 3 * width * height doubles
int launch_raygen(const DCamera& cam, float* d_dirs, double* d_pos, cudaStream_t stream) {
    raygen_walk_kernel<<<(cam.height + 31) / 32, 32, 0, stream>>>(cam, d_pos);
    const size_t n = (size_t)cam.width * cam.height;
    raygen_dir_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(cam, d_pos, d_dirs);
    return (int)cudaGetLastError();
}

// ---- K6: megakernel ------------------------------------------------------------------------------------
// One thread per (pixel, group of sample chunks); a warp covers an 8x4 pixel tile.  The thread walks
// its samples in order and every loop iteration traces exactly ONE ray through the single traversal
// call site — the primary ray, a shadow ray of light `phase`, or the extend ray — so lanes whose path
// ended regenerate immediately and the warp stays converged on the traversal loop.  Per-chunk sums
// are accumulated in sample order in a register and written once: no float atomics, so the image is
// a pure function of (seed, pixel, sample) and of the fixed chunk tree (DESIGN.md "determinism").
enum { PH_PRIMARY = -2, PH_NEWSAMPLE = -1 };

template <bool COUNT, bool REUSE_PRIMARY, bool MIS>
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
        // camera ray through the pixel's top-left corner, precomputed by raygen_kernel (K7)
        const float3 cam_d = f3(__ldg(P.cam_dirs + 3 * (size_t)pixel), __ldg(P.cam_dirs + 3 * (size_t)pixel + 1),
                                __ldg(P.cam_dirs + 3 * (size_t)pixel + 2));
        const float3 cam_o = f3((float)P.cam.eye[0], (float)P.cam.eye[1], (float)P.cam.eye[2]);
        const int nl = sc.n_lights;
        const bool jitter = (P.flags & MCPT_FLAG_JITTER) != 0, anyhit = (P.flags & MCPT_FLAG_ANYHIT_SHADOW) != 0;
        const bool early_out = !(P.flags & MCPT_FLAG_NO_SHADOW_EARLY_OUT);
        float3 prim_d = cam_d;                                          // this sample's primary direction (differs per sample with jitter)
        int chunk = P.chunk_begin + group * P.chunks_per_thread;       // global chunk index
        const int chunk_end = chunk + P.chunks_per_thread;
        int s = (int)((long long)chunk * P.spp / MCPT_NUM_CHUNKS);
        int s_end = (int)((long long)(chunk + 1) * P.spp / MCPT_NUM_CHUNKS);
        float3 sum = f3(0.f, 0.f, 0.f);
        Hit hit0; hit0.k = -1; hit0.t = 0.f; hit0.u = 0.f; hit0.v = 0.f;
        bool have_hit0 = false;
        PathVertex vx; vx.mat = 0; vx.k = 0; vx.t = 0.f; vx.p = vx.pn = vx.kd = vx.wo = f3(0.f, 0.f, 0.f);
        float3 beta = f3(1.f, 1.f, 1.f), L = f3(0.f, 0.f, 0.f), pend = f3(0.f, 0.f, 0.f), wnext = f3(1.f, 1.f, 1.f);
        float3 ro = cam_o, rd = cam_d; float rtmax = 3.0e38f;
        int phase = PH_NEWSAMPLE, lobe = LOBE_DIFFUSE, light_mat = -1;
        MisFrame F; F.n = F.r = F.kd = F.ks = f3(0.f, 0.f, 0.f); F.pd = 1.f; F.Ns = 1.f;
        float pdf_b = 0.f;
        unsigned bounce = 0;
        bool done = false;

        // sample finished with radiance Ls: add in sample order, close chunks as they fill
        auto finish_sample = [&](const float3 Ls) { sum = add_rn(sum, Ls); s++; phase = PH_NEWSAMPLE; };
        // start shading at a (primary or extend) hit that is not a light
        auto begin_path = [&](const Hit& h) {
            if (h.k < 0) { finish_sample(f3(0.f, 0.f, 0.f)); return; }
            const int m = __float_as_int(__ldg(&sc.tri[h.k].v0m).w);
            const int li = sc.materials[m].light_id;
            if (li >= 0) { finish_sample(f3(sc.lights[li].radiance[0], sc.lights[li].radiance[1], sc.lights[li].radiance[2])); return; }
            make_vertex(sc, h, prim_d, vx);
            if (MIS) F = mis_frame(sc, vx);
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
                    else {
                        if (jitter) prim_d = jittered_primary_dir(P.cam, x, y, P.seed_lo, P.seed_hi, pixel, (unsigned)s);
                        phase = PH_PRIMARY; ro = cam_o; rd = prim_d; rtmax = 3.0e38f; have_ray = true;
                    }
                } else if (phase < nl) {
                    const float4 u = philox_uniform4(bounce, STREAM_LIGHT0 + phase, P.seed_lo, P.seed_hi, pixel, (unsigned)s);
                    const DLight& Lt = sc.lights[phase];
                    const bool ok = MIS ? sample_light_mis(sc, Lt, F, vx, u, ro, rd, rtmax, pend) : sample_light(sc, Lt, vx, u, ro, rd, rtmax, pend, anyhit);
                    if (ok) { light_mat = Lt.mat_id; have_ray = true; }
                    else phase++;
                } else {
                    const float4 u = philox_uniform4(bounce, STREAM_PATH, P.seed_lo, P.seed_hi, pixel, (unsigned)s);
                    if (u.x < MCPT_P_RR && bounce + 1 < MCPT_MAX_DEPTH) {   // russian_Roulette (:3-11,:240)
                        bool ext = true;
                        if (MIS) ext = next_ray_mis(sc, F, vx, u, bounce, P.seed_lo, P.seed_hi, pixel, (unsigned)s, ro, rd, lobe, wnext, pdf_b);
                        else next_ray(sc, vx, u, bounce, P.seed_lo, P.seed_hi, pixel, (unsigned)s, ro, rd, lobe, wnext);
                        rtmax = 3.0e38f;
                        if (ext) have_ray = true; else finish_sample(L);
                    } else finish_sample(L);
                }
            }
            if (done) break;
            // ---- the single traversal site ----
            n_rays++;
            Occluder oc = no_occluder();
            if (phase >= 0 && phase < nl) {
                if (anyhit) { oc.t = 3.0e38f; oc.light_mat = -2; }   // any hit settles it
                else if (early_out) oc = shadow_occluder(sc.lights[phase], ro, rd, rtmax);
            }
            const Hit h = trace_closest<COUNT>(sc, ro, rd, rtmax, n_nodes, n_tris, oc);
            // ---- consume the result ----
            if (phase == PH_PRIMARY) {
                if (REUSE_PRIMARY) { hit0 = h; have_hit0 = true; }
                begin_path(h);
            } else if (phase < nl) {
                // visible iff the CLOSEST hit carries the light's material (:212-215)
                const bool visible = anyhit ? h.k < 0 : (h.k >= 0 && __float_as_int(__ldg(&sc.tri[h.k].v0m).w) == light_mat);
                if (visible) L = add_rn(L, mul_rn(beta, pend));
                phase++;
            } else {
                if (h.k < 0) { finish_sample(L); continue; }
                const int m = __float_as_int(__ldg(&sc.tri[h.k].v0m).w);
                const int li = sc.materials[m].light_id;
                const float3 bw = mul_rn(beta, wnext * (1.0f / MCPT_P_RR));
                if (li >= 0) {   // diffuse rays drop emitters (NEE counted them, :247-253); specular/transmission keep them
                    const float3 Le = f3(sc.lights[li].radiance[0], sc.lights[li].radiance[1], sc.lights[li].radiance[2]);
                    if (MIS) L = add_rn(L, mul_rn(bw, Le * mis_weight_bsdf_hit(sc, h.k, rd, h.t, sc.lights[li].area, pdf_b)));
                    else if (lobe != LOBE_DIFFUSE) L = add_rn(L, mul_rn(bw, Le));
                    finish_sample(L);
                } else {
                    beta = bw;
                    make_vertex(sc, h, rd, vx);
                    if (MIS) F = mis_frame(sc, vx);
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
// The record returned through the C ABI is recomputed
// once per ray in FP64 from the uploaded FP64 vertices with the reference's own formulas — plane hit
// (MTPC/sceneManagement.cpp:316-323), t = (p.x - o.x)/d.x (MTPC/pathTracing.cpp:347) and findGarCor's area ratios
// (:416-428) — because FP32 cannot hold t and barycentrics to 1e-5 on the millimetre-sized triangles of the
// cornell-box table seen from 7 units away (|p| * 2^-24 / edge ~ 1e-4).  ~120 FP64 operations per ray against
// ~2000 FP32 instructions of traversal.
__device__ __forceinline__ void refine_hit_fp64(const double* __restrict__ pos, const int face, const float3 of, const float3 df,
                                                float& t_out, float& u_out, float& v_out) {
    const double* q = pos + 9 * (size_t)face;
    const D3 v1 = d3(q[0], q[1], q[2]), v2 = d3(q[3], q[4], q[5]), v3 = d3(q[6], q[7], q[8]);
    const D3 o = d3(of.x, of.y, of.z), d = d3(df.x, df.y, df.z);
    D3 n = dcross(dsub(v1, v2), dsub(v3, v1));                     // Face::calNorm
    const double nl = sqrt(ddot(n, n));
    n = d3(n.x / nl, n.y / nl, n.z / nl);
    const double tp = ddot(dsub(v1, o), n) / ddot(n, d);
    const D3 p = d3(o.x + d.x * tp, o.y + d.y * tp, o.z + d.z * tp);
    t_out = (float)((p.x - o.x) / d.x);
    const D3 e2 = dsub(v1, v3), e3 = dsub(v2, v1);
    const D3 m = dcross(dsub(v3, v2), e2);
    const double an = ddot(m, m);
    u_out = (float)(ddot(dcross(e2, dsub(p, v1)), m) / an);        // b2: weight of v2
    v_out = (float)(ddot(dcross(e3, dsub(p, v2)), m) / an);        // b3: weight of v3
}

// ---- exact hit ids for seam 2 ------------------------------------------------------------------------------
// The integrators decide hits with the FP32 Möller–Trumbore test of trace.cuh; on a ray that runs along an edge shared by two
// triangles both are hit at (nearly) the same t and FP32 rounding picks either.  Seam 2's contract is stricter: the id the
// REFERENCE returns.  K5 therefore uses the FP32 test only as a wide candidate filter (barycentric slack 1e-3, t slack 1e-4) and
// decides every candidate with the reference's own test in FP64, operation for operation and with no FMA contraction (the
// reference is built for baseline x86-64: every product and sum is rounded): the plane hit and inclusive same-side test of
// intersect(Ray&, Face&, Vertex&) (MTPC/sceneManagement.cpp:316-338) with Face::norm from calNorm (:408-412), then
// t = (p.x - o.x) / d.x > 0 and "smallest t wins, ties to the lowest leaf" of bvh_intersect (MTPC/pathTracing.cpp:347-361).
// 1-3 candidates per ray reach the FP64 test; the traversal prunes against the FP32 image of the best t pushed out by 2e-4.
template <bool COUNT>
__device__ __forceinline__ void exact_leaf(const DeviceScene& sc, const int ref, const RayCtx& r, const D3 o64, const D3 d64, Hit& best,
                                           double& best64, unsigned& n_tris) {
    const int k0 = leaf_first(ref), dk = leaf_delta(ref);
    for (int which = 0; which < (dk ? 2 : 1); which++) {
        const int k = which ? k0 + dk : k0;
        const float4 v0 = __ldg(&sc.tri[k].v0m), e1 = __ldg(&sc.tri[k].e1f), e2 = __ldg(&sc.tri[k].e2);
        if (COUNT) n_tris++;
        const float3 E1 = f3(e1), E2 = f3(e2);
        const float3 pvec = cross(r.d, E2);
        const float inv_det = 1.0f / dot(E1, pvec);
        const float3 tvec = r.o - f3(v0);
        const float u = dot(tvec, pvec) * inv_det;
        const float3 qvec = cross(tvec, E1);
        const float v = dot(r.d, qvec) * inv_det;
        const float t = dot(E2, qvec) * inv_det;
        const float slack = 1.0e-3f;
        if (!(u >= -slack && v >= -slack && u + v <= 1.0f + slack && t <= best.t)) continue;   // wide filter; NaN fails it
        double t64;
        if (!reference_tri_test_fp64(sc.pos, sc.leaf_face[k], o64, d64, t64)) continue;
        if (t64 < best64 || (t64 == best64 && k < best.k)) {
            best64 = t64; best.k = k;
            best.t = (float)t64 * 1.0002f + 1.0e-30f;   // pruning bound: never cuts off a candidate that could still tie
        }
    }
}

template <bool COUNT>
__device__ __forceinline__ Hit trace_closest_exact(const DeviceScene& sc, const float3 o, const float3 d, unsigned& n_nodes, unsigned& n_tris) {
    Hit best; best.t = 3.0e38f; best.u = 0.f; best.v = 0.f; best.k = -1;
    if (d.x == 0.0f) return best;   // quirk Q2
    double best64 = 1.0e300;
    const D3 o64 = d3(o.x, o.y, o.z), d64 = d3(d.x, d.y, d.z);
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
        else { exact_leaf<COUNT>(sc, node, r, o64, d64, best, best64, n_tris); node = stack[--sp]; }
    }
    return best;
}

template <bool COUNT>
__global__ void __launch_bounds__(128) trace_kernel(const DeviceScene sc, const float* __restrict__ org_dir, const int n,
                                                    mcpt_hit* __restrict__ hits, unsigned long long* counters) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned n_nodes = 0, n_tris = 0;
    if (i < n) {
        const float* r = org_dir + 6 * (size_t)i;
        const float3 o = f3(r[0], r[1], r[2]), d = f3(r[3], r[4], r[5]);
        const Hit h = trace_closest_exact<COUNT>(sc, o, d, n_nodes, n_tris);
        mcpt_hit out;
        out.t = 0.0f; out.u = 0.0f; out.v = 0.0f; out.leaf = h.k; out.tri = -1;
        if (h.k >= 0) {
            out.tri = sc.leaf_face[h.k];
            refine_hit_fp64(sc.pos, out.tri, o, d, out.t, out.u, out.v);
        }
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
    if (flags & MCPT_FLAG_MIS_BALANCE) pt_megakernel<false, false, true><<<(unsigned)blocks, tpb, 0, stream>>>(P);
    else if (count && reuse) pt_megakernel<true, true, false><<<(unsigned)blocks, tpb, 0, stream>>>(P);
    else if (count) pt_megakernel<true, false, false><<<(unsigned)blocks, tpb, 0, stream>>>(P);
    else if (reuse) pt_megakernel<false, true, false><<<(unsigned)blocks, tpb, 0, stream>>>(P);
    else pt_megakernel<false, false, false><<<(unsigned)blocks, tpb, 0, stream>>>(P);
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
        case 16: sum_planes_kernel<16><<<blocks, tpb, 0, stream>>>(d_planes, count, scale, d_out); break;
        case 32: sum_planes_kernel<32><<<blocks, tpb, 0, stream>>>(d_planes, count, scale, d_out); break;
        case 64: sum_planes_kernel<64><<<blocks, tpb, 0, stream>>>(d_planes, count, scale, d_out); break;
        default: return (int)cudaErrorInvalidValue;
    }
    return (int)cudaGetLastError();
}

}  // namespace mcpt
