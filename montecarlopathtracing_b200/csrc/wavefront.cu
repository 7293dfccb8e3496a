// wavefront.cu — seam 3 as a wavefront integrator (K8 wf_shade + K9 wf_trace), the alternative to the megakernel.
//
// The megakernel's lanes spend most of their time masked off: a warp holds lanes that are shading, drawing Philox
// numbers, walking the BVH and testing triangles all at once (ncu, profiles/r1_megakernel.md: 5.9 of 32 lanes
// active per issued instruction).  Here the two kinds of work run in separate kernels:
//   wf_shade  one thread per path SLOT.  Consumes the slot's ray results, advances the path state machine of
//             shade() (MTPC/pathTracing.cpp:137-266) to the next vertex and emits ALL rays of that vertex at once —
//             one shadow ray per light (NEE) plus the extend ray; their Philox numbers do not depend on each other —
//             into a ray queue compacted with warp ballot/popc and a per-CTA prefix (one atomic per CTA and ray kind).
//   wf_trace  one thread per queued ray: closest hit; extend/primary rays write the hit back to their slot, shadow
//             rays set the slot's visibility bit when the closest hit carries the light's material (:212-215).
// A slot owns (pixel, group of sample chunks) for the whole render, like a megakernel thread: it walks its samples in
// order and adds them up in a register-resident sum that travels in the slot state, so the image is the same pure
// function of (seed, pixel, sample) and of the fixed chunk tree — bit-identical to the megakernel's.
//
// HBM layout (SoA over n_slots; every field a 16-B vector): hit, hit0, sum_s, beta_m, L, rd, pend[n_lights][n_slots];
// vis[n_slots]; ray queue ray_o / ray_d of (n_lights + 1) * n_slots entries.
#include <cstdlib>
#include <vector>
#include <cstdio>
#include "render.h"
#include "shade.cuh"

namespace mcpt {

enum { WF_NEW = 0, WF_WAIT_PRIMARY = 1, WF_WAIT_VERTEX = 2, WF_DONE = 3 };
enum { WF_ID_BITS = 26, WF_MAX_LIGHTS = 32 };   // WF_MAX_PIPES is in render.h
// counters (unsigned, in the scratch block): [0,1] queue tail by iteration parity, [2,3] slots alive by parity,
// [4,5] queue head (dynamic fetch) by parity
enum { WFC_TAIL = 0, WFC_ALIVE = 2, WFC_HEAD = 4, WFC_COUNT = 8 };

struct WfBuffers {
    float4* hit;     // last primary/extend result: (t, u, v, leaf as int bits)
    float4* hit0;    // [pixel]: the pixel's primary hit (quirk Q4: every sample of a pixel shares its primary ray), written by wf_primary
    float4* sum_s;   // chunk sum rgb, next sample index (int bits)
    float4* beta_m;  // throughput rgb, meta bits (see META_*)
    float4* L;       // radiance of the sample in flight
    float4* rd;      // direction of the extend ray in flight (the next vertex's -wo)
    float4* pend;    // [light][slot]: throughput * unshadowed NEE contribution of the vertex in flight
    unsigned* vis;   // shadow-ray results, bit l = light l visible
    unsigned char* warp_done;  // [slot / 32]: every slot of that warp is DONE (or outside the image): wf_shade reads nothing else
    int* next_chunk;           // [pixel]: how many of the launch's chunks the pixel's slots have taken so far (dynamic hand-out)
    float4* ray_o;   // queue: origin, tmax
    float4* ray_d;   // queue: direction, id = slot | (kind << 26), kind 0 = primary/extend, l+1 = shadow ray of light l
    unsigned* counters;
    int n_slots;               // all slots (stride of pend[])
    int slot_begin, slot_end;  // the slots of this pipe
};

// meta: state(2) | lobe(2) << 2 | has_extend << 4 | have_hit0 << 5 | bounce << 8 | chunk << 16
__device__ __forceinline__ unsigned pack_meta(int state, int lobe, bool ext, bool have0, unsigned bounce, int chunk) {
    return (unsigned)state | ((unsigned)lobe << 2) | ((unsigned)ext << 4) | ((unsigned)have0 << 5) | (bounce << 8) | ((unsigned)chunk << 16);
}

// Append a ray for every thread of the CTA with pred set: warp ballots, a shared-memory prefix over the CTA's warps and ONE
// atomic on the queue tail per CTA and call (ncu: with one atomic per warp, 37 % of wf_shade's stall samples sat on the
// return of that same-address atomic).  Rays keep thread order inside the CTA's segment.  Must be called by ALL threads
// of the CTA; `call` alternates the shared scratch so consecutive calls need no extra barrier.
struct PushScratch { unsigned cnt[2][8]; unsigned base[2]; };
__device__ __forceinline__ void push_ray(const bool pred, const WfBuffers& B, unsigned* tail, const float3 o, const float3 d,
                                         const float tmax, const unsigned id, PushScratch& sh, const int call) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, buf = call & 1;
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (lane == 0) sh.cnt[buf][warp] = (unsigned)__popc(mask);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned total = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) { const unsigned c = sh.cnt[buf][w]; sh.cnt[buf][w] = total; total += c; }
        sh.base[buf] = total ? atomicAdd(tail, total) : 0u;
    }
    __syncthreads();
    if (pred) {
        const unsigned i = sh.base[buf] + sh.cnt[buf][warp] + __popc(mask & ((1u << lane) - 1u));
        __stcg(&B.ray_o[i], make_float4(o.x, o.y, o.z, tmax));
        __stcg(&B.ray_d[i], make_float4(d.x, d.y, d.z, __uint_as_float(id)));
    }
}

// ---- K8a: one primary ray per PIXEL ----------------------------------------------------------------------------
// Quirk Q4: all samples of a pixel share one un-jittered primary ray, so its hit is traced once, before the pipes fork, and every
// slot of the pixel starts at the first vertex.  (Until round 2 every SLOT traced it — four times per pixel, 33 M rays at
// 3840 x 2160 — and spent its first two iterations doing so.)  One thread per pixel, a warp per 8 x 4 tile: the rays of a warp are
// as coherent as rays get, which is what the one-ray-per-thread traversal of seam 2 (trace_closest) is good at; it returns the
// same hit as wf_trace, bit for bit (same node step, same triangle test, the closest-hit rule does not depend on visiting order).
template <bool COUNT>
__global__ void __launch_bounds__(128) wf_primary(const RenderParams P, float4* __restrict__ hit0) {
    const int W = P.cam.width, H = P.cam.height;
    const int tiles_x = (W + 7) >> 3, n_tiles = tiles_x * ((H + 3) >> 2);
    const int tile = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    unsigned n_nodes = 0, n_tris = 0;
    if (tile < n_tiles) {
        const int tile_y = (int)fastdiv((unsigned)tile, P.div_tiles_x);
        const int x = (tile - tile_y * tiles_x) * 8 + (lane & 7), y = tile_y * 4 + (lane >> 3);
        if (x < W && y < H) {
            const size_t pixel = (size_t)y * W + x;
            const float3 o = f3((float)P.cam.eye[0], (float)P.cam.eye[1], (float)P.cam.eye[2]);
            const float3 d = f3(__ldg(P.cam_dirs + 3 * pixel), __ldg(P.cam_dirs + 3 * pixel + 1), __ldg(P.cam_dirs + 3 * pixel + 2));
            const Hit h = trace_closest<COUNT>(P.scene, o, d, 3.0e38f, n_nodes, n_tris);
            hit0[pixel] = make_float4(h.t, h.u, h.v, __int_as_float(h.k));
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&P.counters[0], (unsigned long long)W * (unsigned long long)H);
    if (COUNT) {
        const unsigned a = __reduce_add_sync(0xffffffffu, n_nodes), b = __reduce_add_sync(0xffffffffu, n_tris);
        if (lane == 0 && (a | b)) { atomicAdd(&P.counters[1], (unsigned long long)a); atomicAdd(&P.counters[2], (unsigned long long)b); }
    }
}

// ---- K8: shade / path logic ------------------------------------------------------------------------------
#ifndef MCPT_SHADE_MIN_CTAS
#define MCPT_SHADE_MIN_CTAS 4
#endif
template <bool REUSE_PRIMARY, bool MIS>
__global__ void __launch_bounds__(256, MCPT_SHADE_MIN_CTAS) wf_shade(const RenderParams P, const WfBuffers B, const int iter) {
    const DeviceScene& sc = P.scene;
    const int W = P.cam.width, H = P.cam.height;
    const int tiles_x = (W + 7) >> 3, tiles_y = (H + 3) >> 2, n_tiles = tiles_x * tiles_y;
    const int slot = B.slot_begin + blockIdx.x * blockDim.x + threadIdx.x;      // the grid covers whole warps
    const int par = iter & 1;
    unsigned* tail = B.counters + WFC_TAIL + par;
    if (slot == B.slot_begin) {   // the other parity's counters were last used by iteration iter-1, which has completed
        B.counters[WFC_TAIL + (par ^ 1)] = 0u; B.counters[WFC_ALIVE + (par ^ 1)] = 0u; B.counters[WFC_HEAD + (par ^ 1)] = 0u;
    }
    __shared__ PushScratch push_sh;
    const int lane = threadIdx.x & 31;
    const int warp_id = slot >> 5;
    const int group = (int)fastdiv((unsigned)warp_id, P.div_tiles), tile = warp_id - group * n_tiles;
    const int tile_y = (int)fastdiv((unsigned)tile, P.div_tiles_x);
    const int x = (tile - tile_y * tiles_x) * 8 + (lane & 7), y = tile_y * 4 + (lane >> 3);
    const bool valid = slot < B.slot_end && x < W && y < H;
    const unsigned pixel = (unsigned)(y * W + x);
    const int nl = sc.n_lights;
    const bool jitter = (P.flags & MCPT_FLAG_JITTER) != 0, anyhit = (P.flags & MCPT_FLAG_ANYHIT_SHADOW) != 0;

    int state = WF_DONE, lobe = LOBE_DIFFUSE, chunk = 0, s = 0;
    bool has_extend = false, have_hit0 = false, have_vertex = false, emit_primary = false;
    unsigned bounce = 0;
    float3 sum = f3(0.f, 0.f, 0.f), beta = f3(1.f, 1.f, 1.f), L = f3(0.f, 0.f, 0.f), rd = f3(0.f, 0.f, 0.f);
    float3 cam_d = f3(0.f, 0.f, 0.f), vdir = f3(0.f, 0.f, 0.f);
    const float3 cam_o = f3((float)P.cam.eye[0], (float)P.cam.eye[1], (float)P.cam.eye[2]);
    Hit hit0; hit0.t = 0.f; hit0.u = 0.f; hit0.v = 0.f; hit0.k = -1;
    Hit hv = hit0;
    // Chunks are handed out DYNAMICALLY to the slots of a pixel (one atomic counter per pixel): a chunk's plane is the same pure
    // function of (seed, pixel, its samples) whichever slot integrates it — and every slot of a pixel has the same cached
    // primary hit — so the image cannot change, but the slots of a pixel now finish within one chunk of each other.  With the
    // static split a slot whose 64 samples happened to run long kept the render's tail going for ~80 nearly empty iterations.
    const int chunk_end = P.chunk_begin + P.n_chunks;
    (void)group;
    float4 ss_in = make_float4(0.f, 0.f, 0.f, 0.f), h0_in = ss_in, hit_in = ss_in, L_in = ss_in, rd_in = ss_in;
    unsigned vis_in = 0u;
    // Finished warps cost one byte: at 3840 x 2160 the 16:9 frame looks past the box on both sides, ~45 % of the slots are DONE
    // after the first two iterations, and re-reading + re-writing their 16-B meta word every iteration was a quarter of the
    // kernel's DRAM traffic.  A CTA whose eight warps are all finished leaves before the queue barriers.
    bool warp_was_done = false;
    if (iter != 0) {
        unsigned char f = 0;
        if (lane == 0 && slot < B.slot_end) f = __ldcg(&B.warp_done[slot >> 5]);
        warp_was_done = __shfl_sync(0xffffffffu, (int)f, 0) != 0 || (slot - lane) >= B.slot_end;
        if (__syncthreads_and(warp_was_done)) return;
    }
    bool was_done = warp_was_done;

    if (valid && !warp_was_done) {
        cam_d = f3(__ldg(P.cam_dirs + 3 * (size_t)pixel), __ldg(P.cam_dirs + 3 * (size_t)pixel + 1), __ldg(P.cam_dirs + 3 * (size_t)pixel + 2));
        if (REUSE_PRIMARY) h0_in = __ldg(&B.hit0[pixel]);   // wf_primary's result: read-only for the whole loop
        if (iter == 0) {
            state = WF_NEW; chunk = P.chunk_begin + atomicAdd(&B.next_chunk[pixel], 1);
            s = (int)((long long)chunk * P.spp / MCPT_NUM_CHUNKS);
        } else {
            // all of the slot's state is requested at once (independent 16-B loads in flight) and decoded afterwards:
            // the kernel is latency-bound, not bandwidth-bound (ncu: 30 % issue, 21 % DRAM)
            const float4 bm = __ldcg(&B.beta_m[slot]);
            ss_in = __ldcg(&B.sum_s[slot]); hit_in = __ldcg(&B.hit[slot]);
            L_in = __ldcg(&B.L[slot]); rd_in = __ldcg(&B.rd[slot]); vis_in = __ldcg(&B.vis[slot]);
            const unsigned meta = __float_as_uint(bm.w);
            state = meta & 3u; lobe = (meta >> 2) & 3u; has_extend = (meta >> 4) & 1u;
            bounce = (meta >> 8) & 0xffu; chunk = (int)(meta >> 16);
            beta = f3(bm.x, bm.y, bm.z);
            was_done = state == WF_DONE;
            if (state != WF_DONE) {
                sum = f3(ss_in.x, ss_in.y, ss_in.z); s = __float_as_int(ss_in.w);
            }
        }
        if (REUSE_PRIMARY) { hit0.t = h0_in.x; hit0.u = h0_in.y; hit0.v = h0_in.z; hit0.k = __float_as_int(h0_in.w); have_hit0 = true; }
    }
    if (valid && state != WF_DONE) {
        int s_end = (int)((long long)(chunk + 1) * P.spp / MCPT_NUM_CHUNKS);
        bool need_sample = false;
        auto finish_sample = [&](const float3 Ls) { sum = add_rn(sum, Ls); s++; need_sample = true; };
        // first hit of a sample (primary ray, or its cached result)
        auto begin_path = [&](const Hit& h) {
            if (h.k < 0) { finish_sample(f3(0.f, 0.f, 0.f)); return; }
            const int m = __float_as_int(__ldg(&sc.tri[h.k].v0m).w);
            const int li = sc.materials[m].light_id;
            if (li >= 0) { finish_sample(f3(sc.lights[li].radiance[0], sc.lights[li].radiance[1], sc.lights[li].radiance[2])); return; }
            hv = h; have_vertex = true;
            vdir = jitter ? jittered_primary_dir(P.cam, x, y, P.seed_lo, P.seed_hi, pixel, (unsigned)s) : cam_d;
            beta = f3(1.f, 1.f, 1.f); L = f3(0.f, 0.f, 0.f); bounce = 0;
        };
        if (state == WF_NEW) {
            need_sample = true;
        } else if (state == WF_WAIT_PRIMARY) {
            const float4 h4 = hit_in;
            Hit h; h.t = h4.x; h.u = h4.y; h.v = h4.z; h.k = __float_as_int(h4.w);
            begin_path(h);   // only without primary-hit reuse (jitter, MCPT_FLAG_NO_PRIMARY_REUSE): every sample traces its own
        } else {   // WF_WAIT_VERTEX: resolve the vertex's NEE in light order, then look at the extend ray
            const float4 l4 = L_in;
            L = f3(l4.x, l4.y, l4.z);
            const unsigned v = vis_in;
            for (int l = 0; l < nl; l++)
                if ((v >> l) & 1u) { const float4 p4 = __ldcg(&B.pend[(size_t)l * B.n_slots + slot]); L = add_rn(L, f3(p4.x, p4.y, p4.z)); }
            if (!has_extend) {
                finish_sample(L);
            } else {
                const float4 h4 = hit_in;
                Hit h; h.t = h4.x; h.u = h4.y; h.v = h4.z; h.k = __float_as_int(h4.w);
                if (h.k < 0) {
                    finish_sample(L);
                } else {
                    const int m = __float_as_int(__ldg(&sc.tri[h.k].v0m).w);
                    const int li = sc.materials[m].light_id;
                    if (li >= 0) {   // diffuse rays drop emitters (NEE counted them, :247-253); specular/transmission keep them
                        const float3 Le = f3(sc.lights[li].radiance[0], sc.lights[li].radiance[1], sc.lights[li].radiance[2]);
                        if (MIS) {   // balance heuristic against the light's own area sampling (rd.w = pdf of the BSDF sample)
                            const float wb = mis_weight_bsdf_hit(sc, h.k, f3(rd_in.x, rd_in.y, rd_in.z), h.t, sc.lights[li].area, rd_in.w);
                            L = add_rn(L, mul_rn(beta, Le * wb));
                        } else if (lobe != LOBE_DIFFUSE) L = add_rn(L, mul_rn(beta, Le));
                        finish_sample(L);
                    } else {
                        const float4 r4 = rd_in;
                        hv = h; vdir = f3(r4.x, r4.y, r4.z); have_vertex = true; bounce++;
                    }
                }
            }
        }
        while (need_sample) {
            need_sample = false;
            while (s >= s_end && chunk < chunk_end) {   // chunk complete (or empty): write its plane, take the pixel's next one
                float* out = P.planes + (size_t)(chunk - P.chunk_begin) * ((size_t)W * H * 3) + (size_t)pixel * 3;
                out[0] = sum.x; out[1] = sum.y; out[2] = sum.z;
                sum = f3(0.f, 0.f, 0.f);
                chunk = P.chunk_begin + atomicAdd(&B.next_chunk[pixel], 1);
                s = (int)((long long)chunk * P.spp / MCPT_NUM_CHUNKS);
                s_end = (int)((long long)(chunk + 1) * P.spp / MCPT_NUM_CHUNKS);
            }
            if (chunk >= chunk_end) { state = WF_DONE; break; }
            if (REUSE_PRIMARY && have_hit0) begin_path(hit0);
            else emit_primary = true;
        }
        if (have_vertex) state = WF_WAIT_VERTEX;
        else if (emit_primary) state = WF_WAIT_PRIMARY;
    }

    // ---- emission: the whole warp walks the lights together so every push is one ballot + one atomic ----
    PathVertex vx; vx.mat = 0; vx.k = 0; vx.t = 0.f; vx.p = vx.pn = vx.kd = vx.wo = f3(0.f, 0.f, 0.f);
    if (have_vertex) { make_vertex(sc, hv, vdir, vx); __stcg(&B.vis[slot], 0u); }
    MisFrame F;
    if (MIS && have_vertex) F = mis_frame(sc, vx);
    float pdf_b = 0.f;
    for (int l = 0; l < nl; l++) {
        bool ok = false;
        float3 o = f3(0.f, 0.f, 0.f), d = o, contrib = o; float tmax = 0.f;
        if (have_vertex) {
            const float4 u = philox_uniform4(bounce, STREAM_LIGHT0 + l, P.seed_lo, P.seed_hi, pixel, (unsigned)s);
            ok = MIS ? sample_light_mis(sc, sc.lights[l], F, vx, u, o, d, tmax, contrib) : sample_light(sc, sc.lights[l], vx, u, o, d, tmax, contrib, anyhit);
            if (ok) { const float3 bp = mul_rn(beta, contrib); __stcg(&B.pend[(size_t)l * B.n_slots + slot], make_float4(bp.x, bp.y, bp.z, 0.f)); }
        }
        push_ray(ok, B, tail, o, d, tmax, (unsigned)slot | ((unsigned)(l + 1) << WF_ID_BITS), push_sh, l);
    }
    {
        bool ext = false;
        float3 o = cam_o, d = cam_d;
        if (emit_primary && jitter) d = jittered_primary_dir(P.cam, x, y, P.seed_lo, P.seed_hi, pixel, (unsigned)s);
        if (have_vertex) {
            const float4 u = philox_uniform4(bounce, STREAM_PATH, P.seed_lo, P.seed_hi, pixel, (unsigned)s);
            if (u.x < MCPT_P_RR && bounce + 1 < MCPT_MAX_DEPTH) {   // russian_Roulette (:3-11,:240)
                float3 wnext;
                if (MIS) ext = next_ray_mis(sc, F, vx, u, bounce, P.seed_lo, P.seed_hi, pixel, (unsigned)s, o, d, lobe, wnext, pdf_b);
                else { next_ray(sc, vx, u, bounce, P.seed_lo, P.seed_hi, pixel, (unsigned)s, o, d, lobe, wnext); ext = true; }
                if (ext) { beta = mul_rn(beta, wnext * (1.0f / MCPT_P_RR)); rd = d; }
            }
            has_extend = ext;
        }
        push_ray(ext || emit_primary, B, tail, o, d, 3.0e38f, (unsigned)slot, push_sh, nl);
    }
    // ---- store the slot ----
    if (valid && !was_done) {
        __stcg(&B.beta_m[slot], make_float4(beta.x, beta.y, beta.z, __uint_as_float(pack_meta(state, lobe, has_extend, have_hit0, bounce, chunk))));
        if (state != WF_DONE) {
            __stcg(&B.sum_s[slot], make_float4(sum.x, sum.y, sum.z, __int_as_float(s)));
            if (have_vertex) {
                __stcg(&B.L[slot], make_float4(L.x, L.y, L.z, 0.f));
                if (has_extend) __stcg(&B.rd[slot], make_float4(rd.x, rd.y, rd.z, pdf_b));
            }
        }
    } else if (slot < B.slot_end && iter == 0) {
        __stcg(&B.beta_m[slot], make_float4(0.f, 0.f, 0.f, __uint_as_float(pack_meta(WF_DONE, 0, false, false, 0, 0))));
    }
    const bool live = valid && state != WF_DONE;
    if (!warp_was_done) {
        const bool all_done = __all_sync(0xffffffffu, !live);
        if (lane == 0 && slot < B.slot_end && (all_done || iter == 0)) __stcg(&B.warp_done[slot >> 5], (unsigned char)all_done);
    }
    const int alive = __syncthreads_count(live);
    if (threadIdx.x == 0 && alive) atomicAdd(B.counters + WFC_ALIVE + par, (unsigned)alive);
}

// ---- K9: trace the queue -----------------------------------------------------------------------------------
// Persistent warps with per-LANE ray replacement.  Ray lengths are heavy-tailed (a ray that grazes the 15k-triangle
// table visits ten times the nodes of one that hits a wall), so a warp that traces 32 rays to completion idles most
// of its lanes behind the longest ray (ncu: 4.7-6.3 of 32 lanes active).  Here a lane whose ray is finished parks
// its result and, once `refill_min` lanes are parked, the parked lanes write their results back and pull new rays
// from the queue head with one atomic for the warp.  Every loop iteration is one traversal step per lane: lanes at
// an internal node take a node step; lanes at a leaf take a leaf step once `leaf_min` lanes have gathered there (or
// no lane is at a node), so the triangle test is not issued for one or two lanes at a time.
// Two parked leaves per lane instead of one: measured SLOWER in round 1 (cornell-box 1465 -> 1396, veach-mis 680 -> 647 Msamples/s:
// lanes at an internal node per issue rose only 13.6 -> 13.9, the bookkeeping in the node loop cost more).  What did pay (round 2):
// the triangle phase always has TWO test slots (the second belongs to the second triangle of a two-triangle reference); a lane
// whose parked reference holds one triangle and that already stands at its next leaf tests that leaf's first triangle in the free
// slot, and a lane standing at a leaf after the phase parks it there and then, so it walks in the next node step instead of parking
// (lanes per node-step issue on cornell-box 17.1 -> 19.4, node-step issues -12 %, 1 805 -> 1 870 Msamples/s).
#ifndef MCPT_TRACE_MIN_CTAS
#define MCPT_TRACE_MIN_CTAS 9
#endif
template <bool COUNT>
__global__ void __launch_bounds__(128, MCPT_TRACE_MIN_CTAS) wf_trace(const DeviceScene sc, const WfBuffers B, const int iter, unsigned long long* stats,
                                                const int refill_min, const int leaf_min, const int block_min,
                                                const int node_steps, const int flags) {
    const int par = iter & 1;
    const unsigned n = __ldcg(B.counters + WFC_TAIL + par);
    unsigned* head = B.counters + WFC_HEAD + par;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    unsigned n_nodes = 0, n_tris = 0;
    unsigned w_idle = 0, w_fin = 0;
    unsigned w_node = 0, w_tri = 0, w_iter = 0;   // COUNT only: warp-level issues of node steps / triangle tests / loop iterations
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&stats[0], (unsigned long long)n);
    RayCtx r = ray_setup(f3(0.f, 0.f, 0.f), f3(1.f, 1.f, 1.f), sc.qgrid);
    Hit best; best.t = 0.f; best.u = 0.f; best.v = 0.f; best.k = -1;
    extern __shared__ __align__(16) unsigned char wf_smem[];
#if MCPT_SMEM_TOP
    // north_star (b): the top tree levels staged in shared memory — here the first MCPT_SMEM_TOP wide nodes, copied once per CTA
    uint4* top = reinterpret_cast<uint4*>(wf_smem);
    for (int i = threadIdx.x; i < MCPT_SMEM_TOP * 4; i += blockDim.x) {
        const int nd = i >> 2, c = i & 3;
        top[top_slot(nd, c)] = nd < sc.n_wide ? __ldg(&sc.qnodes4[nd].c[c]) : make_uint4(0x0000FFFFu, 0x0000FFFFu, 0x0000FFFFu, 0x80000000u);
    }
    __syncthreads();
#else
    const uint4* top = nullptr;
#endif
#if MCPT_SMEM_STACK
    const SmemStack stack{reinterpret_cast<int*>(wf_smem + MCPT_SMEM_TOP * 64) + threadIdx.x, (int)blockDim.x};
#else
    int stack_mem[MCPT_STACK_SIZE + 4];
    const LocalStack stack{stack_mem};
#endif
    stack.put(0, MCPT_SENT);
    int sp = 1, node = MCPT_SENT;
    unsigned id = 0;
    Occluder oc = no_occluder();
    bool has_ray = false, exhausted = false;
    auto write_back = [&]() {
        const unsigned slot = id & ((1u << WF_ID_BITS) - 1u), kind = id >> WF_ID_BITS;
        if (kind == 0u) {
            __stcg(&B.hit[slot], make_float4(best.t, best.u, best.v, __int_as_float(best.k)));
        } else if (flags & MCPT_FLAG_ANYHIT_SHADOW) {
            if (best.k < 0) atomicOr(&B.vis[slot], 1u << (kind - 1));   // quality extension: visible iff nothing is in between
        } else if (best.k >= 0 && __float_as_int(__ldg(&sc.tri[best.k].v0m).w) == sc.lights[kind - 1].mat_id) {
            atomicOr(&B.vis[slot], 1u << (kind - 1));   // visible iff the CLOSEST hit carries the light's material (:212-215)
        }
    };
    int pend = 0;   // postponed leaf (negative ~k), 0 = none
    while (true) {
        const bool idle = node == MCPT_SENT && pend == 0;
        const unsigned im = __ballot_sync(0xffffffffu, idle);
        if (im != 0u) {
            if (exhausted) {
                if (im == 0xffffffffu) break;
            } else if (__popc(im) >= refill_min || im == 0xffffffffu) {
                if (idle && has_ray) { write_back(); has_ray = false; }
                const int leader = __ffs(im) - 1;
                unsigned base = 0;
                if (lane == leader) base = atomicAdd(head, (unsigned)__popc(im));
                base = __shfl_sync(0xffffffffu, base, leader);
                if (idle) {
                    const unsigned i = base + __popc(im & lt_mask);
                    if (i < n) {
                        const float4 o4 = __ldcg(&B.ray_o[i]), d4 = __ldcg(&B.ray_d[i]);
                        id = __float_as_uint(d4.w);
                        r = ray_setup(f3(o4), f3(d4), sc.qgrid);
                        best.t = o4.w; best.u = 0.f; best.v = 0.f; best.k = -1;
                        sp = 1; has_ray = true;
                        const unsigned kind = id >> WF_ID_BITS;
                        oc = no_occluder();
                        if (kind != 0u) {
                            if (flags & MCPT_FLAG_ANYHIT_SHADOW) { oc.t = 3.0e38f; oc.light_mat = -2; }   // any hit settles it
                            else if (!(flags & MCPT_FLAG_NO_SHADOW_EARLY_OUT)) oc = shadow_occluder(sc.lights[kind - 1], f3(o4), f3(d4), o4.w);
                        }
                        node = (d4.x == 0.0f) ? MCPT_SENT : sc.root;   // quirk Q2: d.x == 0 hits nothing
                    }
                }
                if (base + (unsigned)__popc(im) >= n) exhausted = true;
                continue;
            }
        }
        // a lane that arrives at a leaf parks it and keeps walking: the triangle test is issued when enough lanes hold one
        for (int k = 0; k < node_steps; k++) {
            if (COUNT) {
                const unsigned at_node = __ballot_sync(0xffffffffu, node >= 0);
                const unsigned idle_m = __ballot_sync(0xffffffffu, node == MCPT_SENT && pend == 0);
                const unsigned fin_m = __ballot_sync(0xffffffffu, node == MCPT_SENT && pend != 0);
                if (at_node != 0u && lane == 0) { w_node++; w_idle += __popc(idle_m); w_fin += __popc(fin_m); }
            }
#if MCPT_WIDE
            if (node >= 0) node4_step<COUNT>(sc.qnodes4, node, r, best.t, stack, sp, n_nodes, top);
#else
            if (node >= 0) node_step<COUNT>(sc.qnodes, node, r, best.t, stack_mem, sp, n_nodes);
#endif
            if (node < 0 && node != MCPT_SENT) {
                if (pend == 0) { pend = node; node = stack.get(--sp); }
            }
        }
        const bool has_p = pend != 0;
        // leaf_min <= 1 (the default): the triangle phase runs whenever a lane holds a leaf — measured as good as any
        // threshold (sweeps in DESIGN.md) and it needs no ballots
        unsigned pm = 0xffffffffu, bm = 0u, nm = 0u;
        if (leaf_min > 1 || COUNT) {
            const bool blocked = has_p && node < 0;   // a second leaf, or nothing left but the parked one
            pm = __ballot_sync(0xffffffffu, has_p); bm = __ballot_sync(0xffffffffu, blocked);
            nm = __ballot_sync(0xffffffffu, node >= 0);
        }
        if (COUNT && lane == 0) { w_iter++; if (pm != 0u && (__popc(pm) >= leaf_min || __popc(bm) >= block_min || nm == 0u)) w_tri++; }
        if (has_p && (__popc(pm) >= leaf_min || __popc(bm) >= block_min || nm == 0u)) {
            // Slot 1: the first triangle of the parked reference.  Slot 2: its second triangle; or, when it has none and the lane
            // already stands at another leaf (it can walk no further until this phase), that leaf's first triangle — the lane keeps
            // the rest of that reference, if any, as its node.
            const int k = leaf_first(pend), dk = leaf_delta(pend);
            bool settled = tri_test<COUNT>(sc.tri, k, r, best, n_tris, oc);
            int k2 = dk ? k + dk : -1;
            if (!dk && node < 0 && node != MCPT_SENT) {
                k2 = leaf_first(node);
                const int d2 = leaf_delta(node);
                node = d2 ? ~(k2 + d2) : stack.get(--sp);
            }
            if (k2 >= 0) settled = tri_test<COUNT>(sc.tri, k2, r, best, n_tris, oc) || settled;
            pend = 0;
            if (settled) { node = MCPT_SENT; sp = 1; }   // shadow ray settled: occluded
            // a lane that stands at a leaf parks it here rather than in the next iteration's first node step, which it then walks
            else if (node < 0 && node != MCPT_SENT) { pend = node; node = stack.get(--sp); }
        }
    }
    if (has_ray) write_back();
    if (COUNT) {
        const unsigned a = __reduce_add_sync(0xffffffffu, n_nodes), b = __reduce_add_sync(0xffffffffu, n_tris);
        if (lane == 0) {
            atomicAdd(&stats[1], (unsigned long long)a); atomicAdd(&stats[2], (unsigned long long)b);
            atomicAdd(&stats[3], (unsigned long long)w_node); atomicAdd(&stats[4], (unsigned long long)w_tri); atomicAdd(&stats[5], (unsigned long long)w_iter);
            atomicAdd(&stats[6], (unsigned long long)w_idle); atomicAdd(&stats[7], (unsigned long long)w_fin);
        }
    }
}

// ---- host side ----------------------------------------------------------------------------------------------
static size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }
// developer tuning knobs (defaults are the measured best); read once per render
static int env_int(const char* name, int dflt) { const char* v = getenv(name); return v && *v ? atoi(v) : dflt; }

int wavefront_slots(const RenderParams& P) {
    const long long n_tiles = (long long)((P.cam.width + 7) / 8) * ((P.cam.height + 3) / 4);
    const long long n = n_tiles * 32 * (P.n_chunks / P.chunks_per_thread);
    return n < (1LL << WF_ID_BITS) ? (int)n : -1;
}

size_t wavefront_scratch_bytes(int n_slots, int n_lights, long long n_pixels) {
    const size_t n = (size_t)n_slots;
    return 5 * align256(16 * n) + align256(16 * (size_t)n_pixels) + align256(16 * n * (size_t)(n_lights > 0 ? n_lights : 1)) + align256(4 * n) + align256(n / 32 + 1) + align256(4 * (size_t)n_pixels) +
           2 * align256(16 * n * (size_t)(n_lights + 1)) + align256(4 * WFC_COUNT * WF_MAX_PIPES);
}

void WfHost::destroy() {
    for (cudaStream_t s : side_streams) cudaStreamDestroy(s);
    for (cudaEvent_t e : events) cudaEventDestroy(e);
    for (cudaEvent_t e : trace_events) cudaEventDestroy(e);
    if (h_alive) cudaFreeHost(h_alive);
    side_streams.clear(); events.clear(); trace_events.clear(); h_alive = nullptr;
}

// Runs the shade/trace loop until every slot is done.
//
// The slots are cut into `pipes` contiguous ranges, each with its own ray queue and counters, iterating independently
// on its own stream: while one pipe's wf_trace drains its last long rays (ray lengths are heavy-tailed, so the tail of
// a launch runs at low occupancy) the other pipes' kernels fill the SMs.  The host enqueues `batch` iterations per
// pipe, then reads each pipe's live-slot count of the PREVIOUS batch, so the GPU never waits for the host.
int launch_wavefront(const RenderParams& P, int flags, void* scratch, WfHost& H, cudaStream_t stream, int* launches) {
    const int n_slots = wavefront_slots(P);
    const int nl = P.scene.n_lights;
    if (n_slots <= 0 || nl > WF_MAX_LIGHTS) return (int)cudaErrorInvalidConfiguration;
    cudaError_t e;
#define WF_CU(call) do { e = (call); if (e != cudaSuccess) return (int)e; } while (0)
    int pipes = env_int("MCPT_WF_PIPES", 4);
    pipes = pipes < 1 ? 1 : pipes > WF_MAX_PIPES ? WF_MAX_PIPES : pipes;
    if (n_slots < 256 * 64 * pipes) pipes = 1;   // small images: one pipe already underfills the GPU
    while ((int)H.side_streams.size() < pipes - 1) { cudaStream_t s; WF_CU(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)); H.side_streams.push_back(s); }
    while ((int)H.events.size() < 2 + 3 * WF_MAX_PIPES) { cudaEvent_t ev; WF_CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)); H.events.push_back(ev); }
    if (!H.h_alive) WF_CU(cudaHostAlloc((void**)&H.h_alive, sizeof(unsigned) * 2 * WF_MAX_PIPES, cudaHostAllocDefault));

    WfBuffers B0{};
    char* p = (char*)scratch;
    auto take = [&](size_t bytes) { char* r = p; p += align256(bytes); return r; };
    const size_t n = (size_t)n_slots;
    const size_t n_pixels = (size_t)P.cam.width * P.cam.height;
    B0.hit = (float4*)take(16 * n); B0.hit0 = (float4*)take(16 * n_pixels); B0.sum_s = (float4*)take(16 * n);
    B0.beta_m = (float4*)take(16 * n); B0.L = (float4*)take(16 * n); B0.rd = (float4*)take(16 * n);
    B0.pend = (float4*)take(16 * n * (size_t)(nl > 0 ? nl : 1));
    B0.vis = (unsigned*)take(4 * n);
    B0.warp_done = (unsigned char*)take(n / 32 + 1);
    B0.next_chunk = (int*)take(4 * n_pixels);
    WF_CU(cudaMemsetAsync(B0.next_chunk, 0, 4 * n_pixels, stream));
    B0.ray_o = (float4*)take(16 * n * (size_t)(nl + 1)); B0.ray_d = (float4*)take(16 * n * (size_t)(nl + 1));
    B0.counters = (unsigned*)take(4 * WFC_COUNT * WF_MAX_PIPES);
    B0.n_slots = n_slots;
    WF_CU(cudaMemsetAsync(B0.counters, 0, 4 * WFC_COUNT * WF_MAX_PIPES, stream));

    struct Pipe { WfBuffers B; cudaStream_t st; int shade_blocks; bool done; };
    std::vector<Pipe> pp(pipes);
    const int per = ((n_slots + pipes - 1) / pipes + 255) / 256 * 256;
    for (int k = 0; k < pipes; k++) {
        Pipe& q = pp[k];
        q.B = B0;
        q.B.slot_begin = k * per < n_slots ? k * per : n_slots;
        q.B.slot_end = (k + 1) * per < n_slots ? (k + 1) * per : n_slots;
        q.B.ray_o = B0.ray_o + (size_t)q.B.slot_begin * (nl + 1);
        q.B.ray_d = B0.ray_d + (size_t)q.B.slot_begin * (nl + 1);
        q.B.counters = B0.counters + WFC_COUNT * k;
        q.st = k == 0 ? stream : H.side_streams[k - 1];
        q.shade_blocks = (q.B.slot_end - q.B.slot_begin + 255) / 256;
        q.done = q.shade_blocks == 0;
    }
    const bool count = flags & MCPT_FLAG_COUNT, reuse = !(flags & MCPT_FLAG_NO_PRIMARY_REUSE), mis = (flags & MCPT_FLAG_MIS_BALANCE) != 0;
    int nlaunch = 0;
    if (reuse) {   // the pixels' shared primary hits, once, before the pipes fork
        const unsigned tiles = (unsigned)(((P.cam.width + 7) >> 3) * ((P.cam.height + 3) >> 2));
        if (count) wf_primary<true><<<(tiles + 3) / 4, 128, 0, stream>>>(P, B0.hit0);
        else wf_primary<false><<<(tiles + 3) / 4, 128, 0, stream>>>(P, B0.hit0);
        nlaunch++;
    }
    // fork: the side streams start after everything already enqueued on `stream` (raygen, counter reset, primary hits)
    cudaEvent_t ev_fork = H.events[0];
    WF_CU(cudaEventRecord(ev_fork, stream));
    for (int k = 1; k < pipes; k++) WF_CU(cudaStreamWaitEvent(pp[k].st, ev_fork, 0));

    const int trace_blocks = 148 * env_int("MCPT_TRACE_CTAS_PER_SM", 12) / pipes;   // CTAs of 4 warps, shared by the pipes
    const int refill_min = env_int("MCPT_REFILL_MIN", 12), leaf_min = env_int("MCPT_LEAF_MIN", 1),
              block_min = env_int("MCPT_BLOCK_MIN", 1),
              // wide-node steps between two triangle phases: a ray of a large scene visits more nodes per leaf (5 M triangles: 10.5
              // nodes and 4.0 triangle tests per ray, 3 steps +2.4 %; cornell-box: 6.3 and 4.2, 3 steps -0.6 %)
              node_steps = env_int("MCPT_NODE_STEPS", MCPT_WIDE ? (P.scene.n_faces >= (1 << 20) ? 3 : 2) : 4);
    // A/B builds only (-DMCPT_SMEM_STACK=1 / -DMCPT_SMEM_TOP=K): dynamic shared memory of wf_trace = staged top nodes + a stack of
    // 1 + 3 * (levels of this scene's wide tree) entries per thread, thread-interleaved
    size_t trace_smem = 0;
#if MCPT_SMEM_TOP || MCPT_SMEM_STACK
    trace_smem = (size_t)MCPT_SMEM_TOP * 64 + (MCPT_SMEM_STACK ? (size_t)(4 + 3 * P.scene.wide_levels) * 128 * 4 : 0);
    WF_CU(cudaFuncSetAttribute(wf_trace<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)trace_smem));
    WF_CU(cudaFuncSetAttribute(wf_trace<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)trace_smem));
#endif
    const int batch = 8;
    const bool diag = getenv("MCPT_DIAG_ALIVE") != nullptr;
    int n_timed = 0;
    for (int b = 0;; b++) {
        bool any = false;
        for (int k = 0; k < pipes; k++) {
            Pipe& q = pp[k];
            if (q.done) continue;
            any = true;
            for (int iter = b * batch; iter < (b + 1) * batch; iter++) {
                if (mis) wf_shade<false, true><<<q.shade_blocks, 256, 0, q.st>>>(P, q.B, iter);
                else if (reuse) wf_shade<true, false><<<q.shade_blocks, 256, 0, q.st>>>(P, q.B, iter);
                else wf_shade<false, false><<<q.shade_blocks, 256, 0, q.st>>>(P, q.B, iter);
                while (H.trace_events.size() < 2 * (size_t)(n_timed + 1)) {   // one event pair per wf_trace launch, reused across renders
                    cudaEvent_t ev; WF_CU(cudaEventCreate(&ev));
                    H.trace_events.push_back(ev);
                }
                cudaEventRecord(H.trace_events[2 * n_timed], q.st);
                if (count) wf_trace<true><<<trace_blocks, 128, trace_smem, q.st>>>(P.scene, q.B, iter, P.counters, refill_min, leaf_min, block_min, node_steps, flags);
                else wf_trace<false><<<trace_blocks, 128, trace_smem, q.st>>>(P.scene, q.B, iter, P.counters, refill_min, leaf_min, block_min, node_steps, flags);
                cudaEventRecord(H.trace_events[2 * n_timed + 1], q.st);
                n_timed++; nlaunch += 2;
            }
            const int last = (b + 1) * batch - 1;
            WF_CU(cudaMemcpyAsync(H.h_alive + 2 * k + (b & 1), q.B.counters + WFC_ALIVE + (last & 1), 4, cudaMemcpyDeviceToHost, q.st));
            WF_CU(cudaEventRecord(H.events[2 + 2 * k + (b & 1)], q.st));
        }
        if (!any) break;
        if (b > 0) {   // look at the batch before the one just enqueued
            for (int k = 0; k < pipes; k++) {
                if (pp[k].done) continue;
                WF_CU(cudaEventSynchronize(H.events[2 + 2 * k + ((b - 1) & 1)]));
                if (H.h_alive[2 * k + ((b - 1) & 1)] == 0u) pp[k].done = true;
                if (diag && k == 0) std::fprintf(stderr, "[diag] pipe 0 after iteration %d: %u of %d slots alive\n", b * batch - 1, H.h_alive[2 * k + ((b - 1) & 1)], pp[k].B.slot_end - pp[k].B.slot_begin);
            }
        }
        if (b > (1 << 22)) return (int)cudaErrorUnknown;
    }
    // join: `stream` continues after every pipe has finished
    for (int k = 1; k < pipes; k++) {
        cudaEvent_t ev = H.events[2 + 2 * WF_MAX_PIPES + k];
        WF_CU(cudaEventRecord(ev, pp[k].st));
        WF_CU(cudaStreamWaitEvent(stream, ev, 0));
    }
    H.iterations = n_timed;
    if (launches) *launches += nlaunch;
    return (int)cudaGetLastError();
#undef WF_CU
}

}  // namespace mcpt
