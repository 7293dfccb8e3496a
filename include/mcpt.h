/* mcpt.h — C ABI of the B200-native hot path of Arieys/MonteCarloPathTracing.
 *
 * The reference has no FFI: its hot path sits behind three in-process C++ seams
 * (citations into /root/reference):
 *   seam 1  getMortonCode (MTPC/sceneManagement.cpp:179) + std::sort (MTPC/MTPC.cpp:44)
 *           + BVH::BVH (MTPC/MTPC.cpp:46, MTPC/BVH.cpp:37)          -> mcpt_build_bvh
 *   seam 2  bool ray_intersect(Ray, scene_data&, BVH&, intersection&)
 *           (MTPC/pathTracing.h:21, MTPC/pathTracing.cpp:382)       -> mcpt_trace
 *   seam 3  void generateImg(scene_data&, BVH&, image&, int)
 *           (MTPC/pathTracing.h:25, MTPC/pathTracing.cpp:274)       -> mcpt_render
 * and the outer entry point render_scene(path, filename, spp) (MTPC/MTPC.cpp:35) is kept by the
 * C++ host (montecarlopathtracing_b200/host) and exported here as mcpt_render_scene.
 *
 * Conventions: every function returns 0 on success, a negative mcpt_status otherwise;
 * mcpt_last_error() gives the message (thread-local).  Handles are opaque.  The caller owns all
 * host buffers; the library owns device memory.  One context = one GPU; one host thread per
 * context (mcpt_multi drives several contexts from one call).  Plain pointers and sizes only.  There is NO CPU fallback: without a usable CUDA device
 * every compute call fails with MCPT_ERR_CUDA.
 */
#ifndef MCPT_H
#define MCPT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define MCPT_API __attribute__((visibility("default")))

typedef enum mcpt_status {
    MCPT_OK = 0,
    MCPT_ERR_ARG = -1,     /* bad argument / call order */
    MCPT_ERR_CUDA = -2,    /* CUDA runtime error, no device, or kernels not built for it */
    MCPT_ERR_IO = -3,      /* scene file missing / unparsable */
    MCPT_ERR_NOMEM = -4
} mcpt_status;

/* ---- scene description (host side, flattened from the reference's scene_data) ------------- */

/* Material (MTPC/sceneManagement.h:123-148). tex_id < 0 = no map_Kd. */
typedef struct mcpt_material {
    double kd[3], ks[3];
    double Ns, Ni;
    int32_t tex_id;
    int32_t _pad;
} mcpt_material;

/* Texture: decoded map_Kd, BGR byte order as cv::imread returns it (MTPC/sceneManagement.h:137). */
typedef struct mcpt_texture {
    int32_t height, width;
    const uint8_t* bgr; /* height*width*3 */
} mcpt_texture;

/* Light: one per "mtlname" line of the .camera file, in file order (MTPC/sceneManagement.cpp:241-257). */
typedef struct mcpt_light {
    int32_t mat_id;
    int32_t _pad;
    double radiance[3];
} mcpt_light;

/* Camera (MTPC/sceneManagement.h:150-156). */
typedef struct mcpt_camera {
    double eye[3], lookat[3], up[3];
    double fovy;
    int32_t width, height;
} mcpt_camera;

/* Faces are in OBJ file order (the order of Material::f, i.e. of the light-sampling CDF,
 * MTPC/sceneManagement.cpp:181-182). pos/nrm: 9 doubles per face (v1 v2 v3); uv: 6 per face. */
typedef struct mcpt_scene_desc {
    int32_t n_faces;
    int32_t n_materials;
    int32_t n_textures;
    int32_t n_lights;
    const double* pos;
    const double* nrm;
    const double* uv;
    const int32_t* mat_id;
    const mcpt_material* materials;
    const mcpt_texture* textures;
    const mcpt_light* lights;
} mcpt_scene_desc;

/* Closest-hit record of seam 2. tri = face index in OBJ order (-1 = miss); leaf = position in the
 * Morton-sorted leaf order; (u,v) are the barycentric weights of v2 and v3 (v1 has 1-u-v), i.e.
 * findGarCor's (b2,b3) (MTPC/pathTracing.cpp:416-428). */
typedef struct mcpt_hit {
    float t, u, v;
    int32_t tri;
    int32_t leaf;
} mcpt_hit;

typedef struct mcpt_stats {
    double samples;        /* camera samples integrated */
    double rays;           /* closest-hit queries traced (primary + shadow + extend), device-counted */
    double node_visits;    /* BVH node fetches (only with MCPT_FLAG_COUNT) */
    double tri_tests;      /* ray/triangle tests (only with MCPT_FLAG_COUNT) */
    float render_ms;       /* CUDA-event time of the render kernels on the context's stream */
    float build_ms;        /* last mcpt_build_bvh */
    int32_t kernel_launches; /* kernels launched by the call (render: integrator + reduce) */
    float integrate_ms;    /* CUDA-event time of the integrator alone: the megakernel, or the whole wavefront loop */
    float trace_ms;        /* wavefront mode: summed CUDA-event time of the wf_trace launches (the dominant kernel) */
    int32_t trace_launches;/* wavefront mode: number of wf_trace launches (= loop iterations) */
    int32_t build_launches;/* kernels launched by the last mcpt_build_bvh (ours + CUB's) */
    float exchange_ms;     /* multi-GPU renders, rank 0: CUDA-event time of the one exchange (grouped ncclSend/ncclRecv of the partial sums) */
} mcpt_stats;

enum { MCPT_MODE_MEGAKERNEL = 0, MCPT_MODE_WAVEFRONT = 1 };
enum { MCPT_FLAG_COUNT = 1,            /* count node visits / triangle tests (slower) */
       MCPT_FLAG_NO_PRIMARY_REUSE = 2, /* re-trace the (identical, Q4) primary ray for every sample */
       MCPT_FLAG_NO_SHADOW_EARLY_OUT = 4, /* walk every shadow ray to its closest hit even after its answer is settled (same image, bit for bit) */
       /* Quality extensions.  OFF by default: with either of them the image is no longer the reference's. */
       MCPT_FLAG_JITTER = 8,          /* every sample's primary ray goes through its own Philox-drawn point of the pixel (the reference
                                         shoots all samples through the pixel's corner, MTPC/pathTracing.cpp:305-308): anti-aliasing */
       MCPT_FLAG_ANYHIT_SHADOW = 16,  /* a light sample is visible iff NOTHING lies between the shading point and it (the reference:
                                         iff the closest hit carries the light's material, MTPC/pathTracing.cpp:212-215, so the
                                         far side of a convex light counts as visible through its near side) */
       MCPT_FLAG_MIS_BALANCE = 32     /* a physically based estimator instead of the reference's lobe partition (MTPC/pathTracing.cpp:
                                         171-262): Lambert + normalised Phong on the normalised shading normal, uniform area sampling
                                         of every light and BSDF sampling combined with the balance heuristic, any-hit visibility
                                         (csrc/shade.cuh "MCPT_FLAG_MIS_BALANCE"); implies ANYHIT_SHADOW, no primary-hit reuse */ };

/* Number of sample chunks a render is always cut into (chunk k = samples [k*spp/C, (k+1)*spp/C)); GPU r of R in
 * {1,2,4,8} owns chunks [r*C/R,(r+1)*C/R).  Every chunk is accumulated sequentially into its own plane and the planes are
 * summed by one fixed pairwise tree, so the image depends on neither the GPU count nor on how many chunks one thread (or
 * wavefront slot) integrates — which is what lets the integrator pick its degree of parallelism per workload. */
#define MCPT_NUM_CHUNKS 64

/* ---- host side: the kept C++ shell (loaders, render_scene, PNG) ---------------------------- */

typedef struct mcpt_host_scene mcpt_host_scene;
typedef struct mcpt_ctx mcpt_ctx;

/* Parse <path><name>.{mtl,obj,camera} with the reference's grammar
 * (MTPC/sceneManagement.cpp:17-274; CRLF-safe) and flatten to mcpt_scene_desc.  Textures are read
 * from the "<map_Kd>.bgr" sidecar (int32 rows, int32 cols, BGR bytes), looked up
 * in the CWD, then <path>, then $MCPT_TEXDIR. */
MCPT_API int mcpt_scene_load(const char* path, const char* name, mcpt_host_scene** out);
MCPT_API const mcpt_scene_desc* mcpt_scene_get_desc(const mcpt_host_scene* s);
MCPT_API const mcpt_camera* mcpt_scene_get_camera(const mcpt_host_scene* s);
MCPT_API void mcpt_scene_free(mcpt_host_scene* s);
/* Name of material i as newmtl spelled it (NULL if out of range); valid until mcpt_scene_free. */
MCPT_API const char* mcpt_scene_get_material_name(const mcpt_host_scene* s, int32_t i);

/* render_scene(path, filename, spp) of MTPC/MTPC.cpp:35-68 on cuda device `device`:
 * load -> upload -> build -> render -> quantise -> <out_dir><name>-SPP<spp>.png
 * (out_dir NULL = "../result/", the reference's hard-coded directory, MTPC/MTPC.cpp:64). */
MCPT_API int mcpt_render_scene(const char* path, const char* name, int spp, const char* out_dir, int device);

/* imshow of MTPC/MTPC.cpp:10-33: (uchar)clamp(v*255,0,255), no gamma, uncompressed PNG
 * (replaces MTPC/svpng.inc:77-108; closes the file). rgb = h*w*3 floats, row 0 on top. */
MCPT_API int mcpt_write_png(const char* file, const float* rgb, int width, int height);

/* ---- device side --------------------------------------------------------------------------- */

MCPT_API const char* mcpt_last_error(void);
MCPT_API int mcpt_create(int device, mcpt_ctx** out);
MCPT_API void mcpt_destroy(mcpt_ctx* ctx);
/* Run on the caller's CUDA stream (a cudaStream_t passed as void*) instead of the context's own.  The handle is used as
 * given: NULL is CUDA's legacy default stream (stream 0, what torch.cuda.current_stream().cuda_stream is by default), NOT
 * "back to the context's own stream" — that is mcpt_use_own_stream.  Work the caller enqueues on that stream before / after
 * a call is ordered before / after the call's kernels; calls that return host data synchronise the stream themselves. */
MCPT_API int mcpt_set_stream(mcpt_ctx* ctx, void* cuda_stream);
MCPT_API int mcpt_use_own_stream(mcpt_ctx* ctx);
/* Philox key salt (default 0). Images depend only on (seed, pixel, sample), never on GPU count. */
MCPT_API int mcpt_set_seed(mcpt_ctx* ctx, uint64_t seed);

/* H2D copy of the flattened scene; precomputes per-light area CDFs (MTPC/pathTracing.cpp:177-184). */
MCPT_API int mcpt_upload_scene(mcpt_ctx* ctx, const mcpt_scene_desc* scene);

/* Scene ingest on the GPU: scene_data::read_obj (MTPC/sceneManagement.cpp:76-189) as kernels over
 * the OBJ file's bytes (HOST buffer obj_text, copied H2D once): statements counted and indexed by a scan, numbers converted
 * exactly as atof/strtod rounds them, faces range-checked and flattened straight into the device arrays mcpt_upload_scene
 * would have filled — bit-identical to mcpt_scene_load + mcpt_upload_scene, without the per-face FP64 arrays ever existing
 * on the host.  `tables` supplies materials / textures / lights (its per-face pointers and n_faces are ignored);
 * material_names[n_materials] resolves usemtl.  Errors (MCPT_ERR_IO): non-triangle face, unknown material, index out of
 * range — the first offending line, as the sequential reader meets it — and numbers atof would read but this parser
 * does not (hex floats, inf/nan, > 40 significant digits).  ms_out (may be NULL) = device time incl. the H2D copy. */
MCPT_API int mcpt_ingest_obj(mcpt_ctx* ctx, const char* obj_text, int64_t n_bytes, const mcpt_scene_desc* tables,
                             const char* const* material_names, float* ms_out);
/* The same, reading `file` itself: 8 MiB blocks read by 4 threads into a pinned ring, the H2D copies overlapping the reads. */
MCPT_API int mcpt_ingest_obj_file(mcpt_ctx* ctx, const char* file, const mcpt_scene_desc* tables,
                                  const char* const* material_names, float* ms_out);
/* read_scene (MTPC/sceneManagement.cpp:264-274) with the OBJ on the GPU: <path><name>.mtl and .camera are parsed on the
 * host (a few lines each), the .obj by mcpt_ingest_obj_file.  Leaves the context as mcpt_upload_scene would. */
MCPT_API int mcpt_scene_load_gpu(mcpt_ctx* ctx, const char* path, const char* name, mcpt_camera* camera_out, float* ingest_ms_out);
/* Parity hooks for the ingest: face count of the resident scene (negative = error) and its flattened arrays
 * (9/9/6 doubles and one int32 per face, HOST buffers; any pointer may be NULL). */
MCPT_API int64_t mcpt_scene_face_count(mcpt_ctx* ctx);
MCPT_API int mcpt_get_scene_arrays(mcpt_ctx* ctx, double* pos, double* nrm, double* uv, int32_t* mat_id);

/* Seam 1. Morton codes (bit-exact with MTPC/morton code.cpp:3-32) -> CUB radix sort of (code, face)
 * -> Karras radix tree -> bottom-up AABB refit -> traversal layout.  ms_out (may be NULL) = device time. */
MCPT_API int mcpt_build_bvh(mcpt_ctx* ctx, float* ms_out);

/* Parity hooks for seam 1: n_faces entries each; any pointer may be NULL.
 * codes_obj_order[i] = code of face i; codes_sorted / face_idx_sorted = the sorted leaf order. */
MCPT_API int mcpt_get_morton(mcpt_ctx* ctx, uint32_t* codes_obj_order, uint32_t* codes_sorted,
                             int32_t* face_idx_sorted);
/* The built tree, for structural tests: n_faces-1 internal nodes x 16 floats
 * (see DESIGN.md "node layout"); parent[n_faces-1]; root index. */
MCPT_API int mcpt_get_nodes(mcpt_ctx* ctx, float* nodes16, int32_t* root);
/* The tree traversal walks: the radix tree's leaves re-clustered by surface area (PLOC), same node format
 * and leaf indices; root is n_faces-2.  With MCPT_BVH=lbvh in the environment it is the radix tree itself. */
MCPT_API int mcpt_get_traversal_tree(mcpt_ctx* ctx, float* nodes16, int32_t* root);
/* The 4-wide tree the traversal kernels read: the tree above collapsed breadth-first, root = node 0.
 * nodes16 (may be NULL): capacity max(n_faces-1,1) nodes x 16 uint32 = 4 children x (lo.x | hi.x << 16, lo.y | hi.y << 16,
 * lo.z | hi.z << 16, ref) on the 16-bit grid grid8 = (origin.xyz, 0, cell.xyz, 0): world = origin + q * cell.
 * ref >= 0: wide node; ref < 0: leaf ~(k | d << 26) = sorted triangle k and, when d != 0, k + d; an unused slot has an
 * inverted box (lo = 0xFFFF, hi = 0).  n_wide = number of nodes in use (they are nodes 0 .. n_wide-1). */
MCPT_API int mcpt_get_wide_tree(mcpt_ctx* ctx, uint32_t* nodes16, int32_t* n_wide, float* grid8);

/* Seam 2. Closest hit of n rays, HOST buffers: org_dir = n x 6 floats (origin, direction; the
 * direction need not be unit).  Tie rule and the d.x==0 quirk follow MTPC/pathTracing.cpp:347-361. */
MCPT_API int mcpt_trace(mcpt_ctx* ctx, const float* org_dir, int n, mcpt_hit* hits_out, mcpt_stats* stats);
/* Same with DEVICE buffers (no copies); used to time traversal alone. */
MCPT_API int mcpt_trace_device(mcpt_ctx* ctx, const float* d_org_dir, int n, mcpt_hit* d_hits_out,
                               int flags, mcpt_stats* stats);

/* Seam 3. Mean radiance image into HOST rgb_out (height*width*3 floats, row 0 on top), all
 * MCPT_NUM_CHUNKS sample chunks on this GPU.  width/height override the camera's when > 0. */
MCPT_API int mcpt_render(mcpt_ctx* ctx, const mcpt_camera* cam, int width, int height, int spp, int mode,
                         int flags, float* rgb_out, mcpt_stats* stats);
/* The output stage on the GPU (imshow + svpng, MTPC/MTPC.cpp:10-33, MTPC/svpng.inc:77-108): quantise a DEVICE image of
 * height*width*3 floats and frame it as the same uncompressed PNG mcpt_write_png produces, byte for byte, into the HOST
 * buffer png_out (capacity >= mcpt_png_size(width, height)). */
MCPT_API int64_t mcpt_png_size(int width, int height);
MCPT_API int mcpt_encode_png(mcpt_ctx* ctx, const float* d_rgb, int width, int height, uint8_t* png_out, int64_t capacity);
/* The same stage for callers that must not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * (code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc cannot call
 * mcpt_encode_png, which grows a per-context device buffer).  Needs no mcpt_ctx; calls no cudaMalloc, cudaFree or allocating
 * library routine: the device-side scratch (checksum accumulators, cleared every call, + the encoded file) is one file-scope
 * __device__ array of MCPT_PNG_STATIC_BYTES that every GPU loading the library holds until exit.  Images whose encoded size +
 * 256 B exceeds it (larger than about 1920 x 1080) are rejected with MCPT_ERR_ARG.  d_rgb: DEVICE image on `device`; png_out:
 * host or device buffer of capacity >= mcpt_png_size(); cuda_stream: the caller's stream (NULL = the legacy default stream), which
 * the call synchronises.  Same bytes as mcpt_encode_png.  All calls on a device share the array: calls that overlap on the
 * GPU (other streams or threads) race; serialise them. */
#define MCPT_PNG_STATIC_BYTES (8 << 20)
MCPT_API int mcpt_encode_png_static(int device, const float* d_rgb, int width, int height, uint8_t* png_out, int64_t capacity,
                                    void* cuda_stream);
/* Seam 3 + output stage: render, encode on the device, copy the PNG bytes (not the float image) to the host, write `file`. */
MCPT_API int mcpt_render_png(mcpt_ctx* ctx, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags,
                             const char* file, mcpt_stats* stats);
/* Multi-GPU building block: integrate only chunks [chunk_begin, chunk_end) of the spp samples and
 * leave their fixed-tree SUM (not mean) in the DEVICE buffer d_partial (height*width*3 floats). */
MCPT_API int mcpt_render_chunks(mcpt_ctx* ctx, const mcpt_camera* cam, int width, int height, int spp,
                                int chunk_begin, int chunk_end, int mode, int flags, float* d_partial,
                                mcpt_stats* stats);
/* d_out[i] = scale * pairwise_tree_sum_k(d_partials[k*count + i]), k < n (n a power of two <= 8);
 * the same association as inside mcpt_render, so 1/2/4/8-GPU images are bit-identical. */
MCPT_API int mcpt_sum_partials(mcpt_ctx* ctx, const float* d_partials, int n, int64_t count, float scale,
                               float* d_out);

/* Measurement aid (not on the product path): read bandwidth in GB/s of a 16-B-vector sweep that bypasses L1 over a working set
 * of the given size, repeated `repeats` times — L2 -> SM bandwidth below the L2 size (126 MB on B200), HBM read bandwidth far
 * above it.  bench.py reports the traversal kernels' algorithmic bytes against it. */
MCPT_API int mcpt_measure_read_bandwidth(mcpt_ctx* ctx, int64_t working_set_bytes, int repeats, float* gbs_out);

/* ---- multi-GPU: SPP split, scene replicated, ONE exchange, fixed-order sum ----------------------------
 * The reference's only parallelism is the 8-thread OpenMP loop over one pixel's samples (MTPC/pathTracing.cpp:300-321).
 * Rank r of R in {1,2,4,8} integrates chunks [r*C/R, (r+1)*C/R) of the MCPT_NUM_CHUNKS sample chunks; the R partial sums go to
 * rank 0 in one grouped ncclSend/ncclRecv (straight into one R x count buffer) and are added by the same pairwise tree as inside
 * mcpt_render, so the image does not depend on R, bit for bit.  NCCL (libnccl.so.2) is bound with dlopen on first use: a
 * single-GPU user needs none. */

/* (1) One process per GPU (torchrun, MPI).  Rank 0 makes the id, the launcher broadcasts its MCPT_COMM_ID_BYTES bytes, every
 * rank joins with its own context (ncclCommInitRank).  mcpt_render_dist is COLLECTIVE: every rank calls it with the same
 * arguments after uploading / building the same scene; the mean image arrives in rank 0's HOST buffer rgb_out (ignored on the
 * other ranks; NULL on rank 0 leaves it on the device, see mcpt_device_image).  stats: this rank's samples and rays; rank 0
 * also gets exchange_ms. */
#define MCPT_COMM_ID_BYTES 128
MCPT_API int mcpt_comm_unique_id(char id_out[MCPT_COMM_ID_BYTES]);
MCPT_API int mcpt_comm_init(mcpt_ctx* ctx, int world, int rank, const char id[MCPT_COMM_ID_BYTES]);
MCPT_API int mcpt_comm_destroy(mcpt_ctx* ctx);
MCPT_API int mcpt_render_dist(mcpt_ctx* ctx, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags,
                              float* rgb_out, mcpt_stats* stats);
/* DEVICE pointer of the last mean image mcpt_render / mcpt_render_png / mcpt_render_dist (rank 0) left on this context. */
MCPT_API const float* mcpt_device_image(mcpt_ctx* ctx);

/* (2) One process driving n_dev GPUs (what the C++ host does: render_scene, mcpt_cli --gpus): ncclCommInitAll, one host thread
 * per GPU inside the calls.  devices == NULL means 0 .. n_dev-1.  Same image as one GPU, bit for bit.  stats: samples and rays
 * summed over the GPUs, times = the slowest GPU's. */
typedef struct mcpt_multi mcpt_multi;
MCPT_API int mcpt_multi_create(const int* devices, int n_dev, mcpt_multi** out);
MCPT_API void mcpt_multi_destroy(mcpt_multi* m);
MCPT_API int mcpt_multi_size(const mcpt_multi* m);
MCPT_API mcpt_ctx* mcpt_multi_ctx(mcpt_multi* m, int i);   /* borrowed: GPU i's context (parity hooks) */
MCPT_API int mcpt_multi_set_seed(mcpt_multi* m, uint64_t seed);
MCPT_API int mcpt_multi_upload_scene(mcpt_multi* m, const mcpt_scene_desc* scene);
MCPT_API int mcpt_multi_load_scene_gpu(mcpt_multi* m, const char* path, const char* name, mcpt_camera* camera_out);
MCPT_API int mcpt_multi_build_bvh(mcpt_multi* m, float* ms_out);
MCPT_API int mcpt_multi_render(mcpt_multi* m, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags,
                               float* rgb_out, mcpt_stats* stats);
MCPT_API int mcpt_multi_render_png(mcpt_multi* m, const mcpt_camera* cam, int width, int height, int spp, int mode, int flags,
                                   const char* file, mcpt_stats* stats);
/* render_scene over n_gpus GPUs (0 = every visible GPU, rounded down to a power of two <= 8). */
MCPT_API int mcpt_render_scene_multi(const char* path, const char* name, int spp, const char* out_dir, int n_gpus);

#ifdef __cplusplus
}
#endif
#endif /* MCPT_H */
