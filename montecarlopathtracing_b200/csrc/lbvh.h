// GPU LBVH build (seam 1): buffers owned by a context and the build entry point.
#pragma once
#include "device_types.cuh"

namespace mcpt {

#define W4_SMALL_MAX_N 65536   /* up to this many binary nodes the wide collapse runs as candidates + one cluster (wide.cu) */

struct SceneBuffers {
    int n_faces = 0;
    // uploaded inputs, OBJ order (mcpt_scene_desc)
    double* pos = nullptr; double* nrm = nullptr; double* uv = nullptr; int* mat_id = nullptr;
    // build products
    uint32_t* codes_obj = nullptr;     // Morton code of face i
    uint32_t* codes_sorted = nullptr;  // ascending
    int* iota = nullptr;               // 0..n-1 (sort values in)
    int* leaf_face = nullptr;          // sorted leaf k -> OBJ face index
    int* leaf_parent = nullptr;
    unsigned* flags = nullptr;         // refit arrival counters, one per internal node
    BvhNode* nodes = nullptr;          // max(n-1,1) internal nodes; root = node 0
    TriGeom* tri = nullptr;            // sorted order
    TriShade* shade = nullptr;         // sorted order
    float* root_box = nullptr;         // 6 floats: lo.xyz, hi.xyz
    QNode* qnodes = nullptr;           // traversal nodes, same indexing as `nodes`
    QGrid* qgrid = nullptr;
    QNode4* qnodes4 = nullptr;         // 4-wide traversal nodes (wide.cu), breadth-first, root = 0
    int* w4_queue = nullptr;           // wide node -> the binary node it is rooted at
    QNode4* w4_cand = nullptr;         // small scenes (<= W4_SMALL_MAX_N nodes): every binary node as the root of a wide node (wide.cu)
    unsigned* w4_ctr = nullptr;        // 16 words: queue tail (= number of wide nodes), processed, finished, error, internal levels, K1's bad-vertex flag, -, -, three per-level counters
    int w4_resident_blocks = 0;        // co-resident CTAs of wide_collapse_kernel on this context's device (occupancy query)
    unsigned* build_err = nullptr;     // device flag set by K1 when a vertex is not finite or beyond +-1e18 (checked by mcpt_build_bvh)
    void* cub_temp = nullptr; size_t cub_temp_bytes = 0;
    // quality pass (ploc.cu): the tree traversal walks
    bool use_ploc = true;              // MCPT_BVH=lbvh keeps the radix tree for traversal (A/B measurements)
    bool fold_pairs = true;            // MCPT_LEAF_PAIRS=0 keeps one triangle per leaf reference (A/B measurements)
    BvhNode* nodes2 = nullptr;         // PLOC tree, n-1 nodes, root = n-2
    float4* cl_lo[2] = {nullptr, nullptr}; float4* cl_hi[2] = {nullptr, nullptr};   // cluster boxes (+ref in lo.w), ping-pong
    unsigned char* ploc_flag = nullptr;   // one byte per cluster: survives | merges-as-left << 1 | (partner offset + RADIUS) << 2
    void* ploc_temp = nullptr; size_t ploc_temp_bytes = 0; void* ploc_state = nullptr;   // per-CTA counts; PlocState
    int ploc_resident_blocks = 0;      // co-resident CTAs of ploc_rounds_kernel on this context's device
    bool walk_is_radix = false;        // the traversal tree is the radix tree (MCPT_BVH=lbvh, or the quality pass fell back)
    int root = 0;                      // root of the traversal tree (0: radix tree, n-2: PLOC tree)
    int n_wide = 0, wide_levels = 0;   // wide nodes in use / internal levels of the wide tree (read back by mcpt_build_bvh)
    void* arena = nullptr;             // every pointer above is carved from this one allocation (api.cu alloc_scene)
    bool built = false;
};

// Bytes of CUB temp storage needed for n faces.
size_t lbvh_sort_temp_bytes(int n);
// K1 Morton -> CUB radix sort -> K2 radix tree -> K3 leaf emit + atomic bottom-up refit -> K4 node quantisation.
// Returns a cudaError_t as int; all work is enqueued on `stream`.
// side: a second stream of the context, ev_fork / ev_join: timing-disabled events (the build forks after the sort, see lbvh.cu).
struct BuildStreams { cudaStream_t side; cudaEvent_t ev_fork, ev_join; };
int lbvh_build(SceneBuffers& sb, cudaStream_t stream, const BuildStreams& bs, int* launches);
// Redo the wide collapse over the radix tree (quality pass stalled / too deep); enqueued on `stream`.
int lbvh_fallback_to_radix(SceneBuffers& sb, cudaStream_t stream, int* launches);
// K4a (ploc.cu): parallel locally-ordered clustering over the Morton-ordered leaf boxes -> sb.nodes2.
size_t ploc_scratch_bytes(int n);
int ploc_build(SceneBuffers& sb, cudaStream_t stream, int* launches, bool init_w4);
// After the stream has been synchronised: did the quality pass finish (false: stalled, sb.nodes2 is unusable)?
bool ploc_ok(const SceneBuffers& sb, cudaError_t* err);

// K4b (wide.cu): collapse the binary traversal tree `walk` (root `root2`) into sb.qnodes4.
// prepared: the queue is already filled with W4_EMPTY and the counters are reset (K3 and the PLOC finish kernel do that on the way).
int wide_build(SceneBuffers& sb, const BvhNode* walk, int root2, cudaStream_t stream, int* launches, bool prepared);
// Internal levels of the wide tree a traversal stack of MCPT_STACK_SIZE entries can serve (3 pushes per level + the sentinel).
#define MCPT_MAX_WIDE_LEVELS 21

}  // namespace mcpt
