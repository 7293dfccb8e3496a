// obj_ingest.h — GPU-side scene ingest: scene_data::read_obj (MTPC/sceneManagement.cpp:76-189)
// as four kernels over the OBJ text in device memory.  See obj_ingest.cu.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <string>
#include <vector>

namespace mcpt {

enum ObjErrorCode { OBJ_ERR_NOT_TRIANGLE = 1, OBJ_ERR_UNKNOWN_MATERIAL = 2, OBJ_ERR_INDEX_RANGE = 3, OBJ_ERR_NUMBER = 4 };

// Grow-only device scratch owned by the context: the ingest's temporaries are carved from two of these (one sized from the
// text length, one from the statement counts), so a second ingest allocates nothing.
struct DevArena {
    char* base = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap && base) return cudaSuccess;
        if (base) cudaFree(base);
        base = nullptr; cap = 0;
        cudaError_t e = cudaMalloc((void**)&base, bytes ? bytes : 1);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() { if (base) cudaFree(base); base = nullptr; cap = 0; }
};

struct ObjIngest {
    DevArena* arena_text = nullptr;   // counts, scan scratch, defer flags, status
    DevArena* arena_stmt = nullptr;   // parsed statements
    const char* d_text = nullptr;     // the OBJ file, bytes as on disk (not owned)
    unsigned n_bytes = 0;
    unsigned n_chunks = 0;
    unsigned n_v = 0, n_vn = 0, n_vt = 0, n_f = 0, n_um = 0;   // statement counts (host copies)
    // device temporaries (carved from the arenas)
    void* counts = nullptr;           // per-chunk statement counts, then their exclusive scan
    void* scan_temp = nullptr; size_t scan_temp_bytes = 0;
    unsigned char* defer = nullptr;   // per chunk: holds a number that needs the 896-bit conversion
    double *v = nullptr, *vn = nullptr, *vt = nullptr;
    int* fidx = nullptr;              // 9 per face: (v, vn, vt) x 3 corners, 0-based; fidx[9f] < 0 = rejected face
    unsigned* f_um = nullptr;         // usemtl statements before the face
    unsigned* f_off = nullptr;        // byte offset of the face's line (error reports)
    unsigned* um_off = nullptr;       // per usemtl statement: offset / length of the name token
    unsigned* um_len = nullptr;
    int* um_mat = nullptr;            // per usemtl statement: material index (host-resolved), -1 = unknown
    unsigned long long* status = nullptr;   // [0] = min over errors of (line offset << 8 | code), [1] = deferred chunks
    unsigned long long h_status[2] = {~0ull, 0ull};
    int launches = 0;
};

// Stage 1: classify + count (K16), scan (CUB), parse (K17; K17b for deferred chunks).  Synchronises the stream; fills the
// counts and h_status.  Returns a cudaError_t as int.
int obj_ingest_parse(ObjIngest& g, cudaStream_t stream);
// The names of the usemtl statements, in file order (D2H of a packed buffer).
int obj_ingest_names(ObjIngest& g, std::vector<std::string>& names, cudaStream_t stream);
// Stage 2: per-face material + gather into the flattened scene arrays (K18).  um_mat: host array of n_um material ids.
// Synchronises the stream and refreshes h_status.
int obj_ingest_flatten(ObjIngest& g, const int* um_mat, double* d_pos, double* d_nrm, double* d_uv, int* d_mat_id, cudaStream_t stream);
// out[18 j .. 18 j + 17] = pos[9 idx[j] ..], nrm[9 idx[j] ..]
int launch_gather_faces(const int* d_idx, int m, const double* d_pos, const double* d_nrm, double* d_out, cudaStream_t stream);

}  // namespace mcpt
