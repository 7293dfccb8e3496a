// obj_ingest.cu — scene ingest on the GPU: the OBJ reader of the reference
// (scene_data::read_obj, MTPC/sceneManagement.cpp:76-189: getline, substr/find tokenising, atof/atoi, one push_back per
// statement, a copy of three vertices/normals/texcoords into every Face) restated as data-parallel passes over the
// file's bytes in device memory.  Grammar = host/scene_loader.cpp's (the reference's, made CRLF- and blank-tolerant),
// and the result is bit-identical to it: same doubles (obj_number.cuh), same face order, same errors.
//
//   K16 obj_count_kernel    one thread per 128-byte chunk of text: the statements (v / vn / vt / f / usemtl) whose line
//                           STARTS in the chunk are classified by their first token and counted
//        cub ExclusiveScan  of the per-chunk counts: every statement's index inside its array (file order)
//   K17 obj_parse_kernel    same walk, now parsing: v/vn/vt -> FP64 arrays; f -> 9 indices, range-checked against the
//                           number of v/vn/vt statements BEFORE the line (what indexing a growing std::vector checks);
//                           usemtl -> offset/length of the name.  Numbers that need the 896-bit conversion mark their
//                           chunk, and K17b (the same kernel, BIG = true) redoes only marked chunks.
//   K18 obj_flatten_kernel  one thread per face corner: material of the last usemtl before the face, then the gather of
//                           position / normal / texcoord into the per-face FP64 arrays the LBVH build and the hit
//                           refinement read (DeviceScene::pos etc.)
// Errors are reported as the minimum over (line offset << 8 | code): the first offending line of the file, as a
// sequential reader would meet it.
#include <cub/device/device_scan.cuh>
#include "obj_ingest.h"
#include "obj_number.cuh"

namespace mcpt {

#define OBJ_CHUNK 128u
enum { L_NONE = 0, L_V, L_VN, L_VT, L_F, L_UM };

struct __align__(16) ObjCounts { unsigned v, vn, vt, f, um, pad0, pad1, pad2; };
struct ObjCountsSum {
    __host__ __device__ ObjCounts operator()(const ObjCounts& a, const ObjCounts& b) const {
        ObjCounts r;
        r.v = a.v + b.v; r.vn = a.vn + b.vn; r.vt = a.vt + b.vt; r.f = a.f + b.f; r.um = a.um + b.um;
        r.pad0 = r.pad1 = r.pad2 = 0;
        return r;
    }
};

__device__ __forceinline__ bool blank(char c) { return c == ' ' || c == '\t' || c == '\r'; }

// The block's stretch of text staged in shared memory.  A thread walks its own 128-byte chunk byte by byte, so a warp's
// 32 loads are 128 bytes apart: from global memory that is 32 L1 wavefronts per load instruction (what bounded the first
// version of these kernels).  In shared memory one pad word per 128 bytes puts the 32 chunks on 32 different banks.
// Bytes outside the staged window (a line longer than the look-ahead) fall back to global memory.
#define OBJ_BLOCK 128u
#define OBJ_SPAN (OBJ_BLOCK * OBJ_CHUNK)   // 16 KiB of text per block
#define OBJ_BACK 16u                       // look-behind (one byte is needed; 16 keeps the window aligned)
#define OBJ_LOOK 2048u                     // look-ahead for lines that run past the block's end
#define OBJ_SMEM_WORDS (((OBJ_BACK + OBJ_SPAN + OBJ_LOOK) / 4u) * 33u / 32u + 2u)

struct TextView {
    const char* g; const unsigned char* s; unsigned lo, len;
    __device__ __forceinline__ char operator[](unsigned i) const {
        const unsigned x = i - lo;
        return x < len ? (char)s[x + ((x >> 7) << 2)] : g[i];
    }
};
struct TextIt {                            // the iterator obj_number.cuh parses through
    const TextView* v; unsigned i;
    __device__ __forceinline__ char operator*() const { return (*v)[i]; }
    __device__ __forceinline__ char operator[](int k) const { return (*v)[i + k]; }
    __device__ __forceinline__ TextIt& operator++() { i++; return *this; }
    __device__ __forceinline__ TextIt operator++(int) { TextIt t = *this; i++; return t; }
    __device__ __forceinline__ TextIt operator+(int k) const { return TextIt{v, i + k}; }
    __device__ __forceinline__ int operator-(const TextIt& o) const { return (int)(i - o.i); }
    __device__ __forceinline__ bool operator<(const TextIt& o) const { return i < o.i; }
    __device__ __forceinline__ bool operator==(const TextIt& o) const { return i == o.i; }
};

// Stage [block_begin - OBJ_BACK, block_begin + OBJ_SPAN + OBJ_LOOK) ∩ [0, n); the text buffer has 16 spare bytes.
__device__ __forceinline__ TextView stage_text(const char* __restrict__ t, unsigned n, unsigned* smem) {
    const unsigned b0 = blockIdx.x * OBJ_SPAN;
    TextView v;
    v.g = t; v.s = (const unsigned char*)smem;
    v.lo = b0 >= OBJ_BACK ? b0 - OBJ_BACK : 0u;
    const unsigned long long want = (unsigned long long)b0 + OBJ_SPAN + OBJ_LOOK;
    const unsigned hi = want < n ? (unsigned)want : n;
    v.len = hi > v.lo ? hi - v.lo : 0u;
    const unsigned* src = (const unsigned*)(t + v.lo);
    for (unsigned w = threadIdx.x; w < (v.len + 3u) / 4u; w += blockDim.x) smem[w + (w >> 5)] = __ldg(src + w);
    __syncthreads();
    return v;
}

// first line start inside [c0, c1), or 0xFFFFFFFF
__device__ __forceinline__ unsigned first_line_start(const TextView& t, unsigned c0, unsigned c1, unsigned n) {
    if (c0 == 0) return 0;
    const unsigned lim = c1 < n ? c1 : n;
    unsigned j = c0 - 1;
    while (j < lim && t[j] != '\n') j++;
    return j < lim ? j + 1 : 0xFFFFFFFFu;
}

// The line [s, le): kind from its first token; `rest` = first byte after that token (for usemtl: start of the name)
__device__ __forceinline__ int classify_line(const TextView& t, unsigned s, unsigned n, unsigned& le, unsigned& rest) {
    unsigned p = s;
    while (p < n && blank(t[p])) p++;                   // '\n' is not a blank: stops at the line end
    unsigned q = p;
    char c0 = 0, c1 = 0;
    bool usemtl = true;
    while (q < n) {
        const char c = t[q];
        if (c == '\n' || blank(c)) break;
        const unsigned k = q - p;
        if (k == 0) c0 = c; else if (k == 1) c1 = c;
        if (k < 6) usemtl = usemtl && c == "usemtl"[k];
        q++;
    }
    rest = q;
    unsigned e = q;
    bool more = false;                                  // a second token on the line
    unsigned r = q;
    while (e < n) {
        const char c = t[e];
        if (c == '\n') break;
        if (!more && !blank(c)) { more = true; r = e; }
        e++;
    }
    le = e;
    if (e - s < 2) return L_NONE;                       // scene_loader.cpp: lines shorter than 2 bytes are skipped
    const unsigned len = q - p;
    if (len == 1) return c0 == 'v' ? L_V : (c0 == 'f' ? L_F : L_NONE);
    if (len == 2 && c0 == 'v') return c1 == 'n' ? L_VN : (c1 == 't' ? L_VT : L_NONE);
    if (len == 6 && usemtl && more) { rest = r; return L_UM; }   // a bare "usemtl" changes nothing
    return L_NONE;
}

__global__ void __launch_bounds__(128) obj_count_kernel(const char* __restrict__ text, unsigned n, unsigned n_chunks, ObjCounts* __restrict__ counts) {
    __shared__ unsigned smem[OBJ_SMEM_WORDS];
    const TextView t = stage_text(text, n, smem);
    const unsigned c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c > n_chunks) return;
    ObjCounts k = {0, 0, 0, 0, 0, 0, 0, 0};
    if (c < n_chunks) {                                 // entry n_chunks stays zero: its scan value is the total
        const unsigned c0 = c * OBJ_CHUNK, c1 = c0 + OBJ_CHUNK;
        unsigned s = first_line_start(t, c0, c1, n);
        while (s < c1 && s < n) {
            unsigned le, rest;
            const int kind = classify_line(t, s, n, le, rest);
            k.v += kind == L_V; k.vn += kind == L_VN; k.vt += kind == L_VT; k.f += kind == L_F; k.um += kind == L_UM;
            s = le + 1;
        }
    }
    counts[c] = k;
}

__device__ __forceinline__ bool next_tok(const TextView& t, unsigned& p, unsigned e, unsigned& b, unsigned& en) {
    while (p < e && blank(t[p])) p++;
    if (p >= e) return false;
    b = p;
    while (p < e && !blank(t[p])) p++;
    en = p;
    return true;
}
__device__ __forceinline__ void report(unsigned long long* status, unsigned line, int code) {
    atomicMin(&status[0], ((unsigned long long)line << 8) | (unsigned long long)code);
}

// "a/b/c" -> 0-based (v, vn, vt); missing parts repeat the first (scene_loader.cpp parse_corner)
__device__ __forceinline__ void parse_corner(const TextView& t, unsigned b, unsigned en, long long idx[3]) {
    const TextIt end{&t, en};
    idx[0] = objnum::parse_long(TextIt{&t, b}, end) - 1;
    idx[1] = idx[2] = idx[0];
    unsigned s1 = b;
    while (s1 < en && t[s1] != '/') s1++;
    if (s1 == en) return;
    idx[1] = objnum::parse_long(TextIt{&t, s1 + 1}, end) - 1;
    unsigned s2 = s1 + 1;
    while (s2 < en && t[s2] != '/') s2++;
    if (s2 == en) return;
    idx[2] = objnum::parse_long(TextIt{&t, s2 + 1}, end) - 1;
}

template <bool BIG>
__device__ __forceinline__ void parse_numbers(const TextView& t, unsigned p, unsigned le, unsigned line, int count, double* __restrict__ out,
                                              bool& deferred, unsigned long long* status) {
    for (int k = 0; k < count; k++) {
        unsigned b, en;
        double val = 0.0;                               // a missing token reads as 0 (Tokens::num)
        if (next_tok(t, p, le, b, en)) {
            const int st = objnum::parse_double<BIG>(TextIt{&t, b}, TextIt{&t, en}, val);
            if (st == objnum::NUM_DEFER) deferred = true;
            else if (st == objnum::NUM_UNSUPPORTED) report(status, line, OBJ_ERR_NUMBER);
        }
        out[k] = val;
    }
}

template <bool BIG>
__global__ void __launch_bounds__(128) obj_parse_kernel(const char* __restrict__ text, unsigned n, unsigned n_chunks, const ObjCounts* __restrict__ base,
                                                        unsigned char* __restrict__ defer, double* __restrict__ v, double* __restrict__ vn,
                                                        double* __restrict__ vt, int* __restrict__ fidx, unsigned* __restrict__ f_um,
                                                        unsigned* __restrict__ f_off, unsigned* __restrict__ um_off, unsigned* __restrict__ um_len,
                                                        unsigned long long* __restrict__ status) {
    __shared__ unsigned smem[OBJ_SMEM_WORDS];
    const unsigned c = blockIdx.x * blockDim.x + threadIdx.x;
    const bool mine = c < n_chunks && (!BIG || defer[c]);
    if (BIG && !__syncthreads_or(mine)) return;         // no chunk of this block asked for the 896-bit pass
    const TextView t = stage_text(text, n, smem);
    if (!mine) return;
    ObjCounts k = base[c];
    const unsigned c0 = c * OBJ_CHUNK, c1 = c0 + OBJ_CHUNK;
    unsigned s = first_line_start(t, c0, c1, n);
    bool deferred = false;
    while (s < c1 && s < n) {
        unsigned le, rest;
        const int kind = classify_line(t, s, n, le, rest);
        if (kind == L_V) { parse_numbers<BIG>(t, rest, le, s, 3, v + 3ull * k.v, deferred, status); k.v++; }
        else if (kind == L_VN) { parse_numbers<BIG>(t, rest, le, s, 3, vn + 3ull * k.vn, deferred, status); k.vn++; }
        else if (kind == L_VT) { parse_numbers<BIG>(t, rest, le, s, 2, vt + 2ull * k.vt, deferred, status); k.vt++; }
        else if (kind == L_UM) {
            unsigned q = rest;
            while (q < le && !blank(t[q])) q++;
            um_off[k.um] = rest; um_len[k.um] = q - rest;
            k.um++;
        } else if (kind == L_F && !BIG) {
            unsigned p = rest, b[4], en[4];
            int ntok = 0;
            while (ntok < 4 && next_tok(t, p, le, b[ntok], en[ntok])) ntok++;
            int* o = fidx + 9ull * k.f;
            bool ok = ntok == 3;                        // "f" + exactly three corners
            if (!ok) report(status, s, OBJ_ERR_NOT_TRIANGLE);
            else {
                for (int cnr = 0; cnr < 3; cnr++) {
                    long long idx[3];
                    parse_corner(t, b[cnr], en[cnr], idx);
                    if (idx[0] < 0 || idx[0] >= (long long)k.v || idx[1] < 0 || idx[1] >= (long long)k.vn || idx[2] < 0 || idx[2] >= (long long)k.vt) ok = false;
                    else { o[3 * cnr] = (int)idx[0]; o[3 * cnr + 1] = (int)idx[1]; o[3 * cnr + 2] = (int)idx[2]; }
                }
                if (!ok) report(status, s, OBJ_ERR_INDEX_RANGE);
            }
            if (!ok) o[0] = -1;
            f_um[k.f] = k.um; f_off[k.f] = s;
            k.f++;
        }
        s = le + 1;
    }
    if (!BIG && deferred) { defer[c] = 1; atomicAdd(&status[1], 1ull); }
}

__global__ void __launch_bounds__(256) obj_flatten_kernel(unsigned n_f, const int* __restrict__ fidx, const unsigned* __restrict__ f_um,
                                                          const unsigned* __restrict__ f_off, const int* __restrict__ um_mat,
                                                          const double* __restrict__ v, const double* __restrict__ vn, const double* __restrict__ vt,
                                                          double* __restrict__ pos, double* __restrict__ nrm, double* __restrict__ uv,
                                                          int* __restrict__ mat_id, unsigned long long* __restrict__ status) {
    const unsigned long long g = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= 3ull * n_f) return;
    const unsigned f = (unsigned)(g / 3ull), cnr = (unsigned)(g - 3ull * f);
    const int* o = fidx + 9ull * f;
    if (cnr == 0) {
        const unsigned um = f_um[f];
        const int m = um ? um_mat[um - 1] : -1;
        if (m < 0) report(status, f_off[f], OBJ_ERR_UNKNOWN_MATERIAL);
        mat_id[f] = m < 0 ? 0 : m;
    }
    if (o[0] < 0) return;
    const double* a = v + 3ull * o[3 * cnr]; const double* b = vn + 3ull * o[3 * cnr + 1]; const double* c = vt + 2ull * o[3 * cnr + 2];
    double* P = pos + 9ull * f + 3 * cnr; double* N = nrm + 9ull * f + 3 * cnr; double* U = uv + 6ull * f + 2 * cnr;
    P[0] = a[0]; P[1] = a[1]; P[2] = a[2];
    N[0] = b[0]; N[1] = b[1]; N[2] = b[2];
    U[0] = c[0]; U[1] = c[1];
}

__global__ void obj_pack_names_kernel(const char* __restrict__ t, unsigned n_um, const unsigned* __restrict__ um_off, const unsigned* __restrict__ um_len,
                                      const unsigned* __restrict__ dst, char* __restrict__ out) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_um) return;
    for (unsigned k = 0; k < um_len[i]; k++) out[dst[i] + k] = t[um_off[i] + k];
}

__global__ void gather_faces_kernel(const int* __restrict__ idx, int m, const double* __restrict__ pos, const double* __restrict__ nrm, double* __restrict__ out) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= m * 18) return;
    const int j = g / 18, k = g - 18 * j;
    out[g] = k < 9 ? pos[9ull * idx[j] + k] : nrm[9ull * idx[j] + (k - 9)];
}

namespace {
#define OBJ_CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return (int)e_; } while (0)
struct Carver {
    char* base; size_t used = 0;
    explicit Carver(char* b) : base(b) {}
    template <class T> void take(T*& p, size_t count) {
        if (base) p = (T*)(base + used);
        used += (sizeof(T) * (count ? count : 1) + 255) & ~(size_t)255;
    }
};
}  // namespace

int obj_ingest_parse(ObjIngest& g, cudaStream_t st) {
    const unsigned n = g.n_bytes;
    g.n_chunks = (n + OBJ_CHUNK - 1) / OBJ_CHUNK;
    const unsigned nc1 = g.n_chunks + 1;
    ObjCounts* counts = nullptr;
    ObjCounts zero = {0, 0, 0, 0, 0, 0, 0, 0};
    OBJ_CU(cub::DeviceScan::ExclusiveScan(nullptr, g.scan_temp_bytes, counts, counts, ObjCountsSum(), zero, (int)nc1, st));
    char* scan_temp = nullptr;
    for (int pass = 0; pass < 2; pass++) {
        Carver c(pass ? g.arena_text->base : nullptr);
        c.take(counts, nc1); c.take(g.defer, g.n_chunks); c.take(g.status, 2); c.take(scan_temp, g.scan_temp_bytes);
        if (pass == 0) OBJ_CU(g.arena_text->reserve(c.used));
    }
    g.counts = counts; g.scan_temp = scan_temp;
    OBJ_CU(cudaMemsetAsync(g.defer, 0, g.n_chunks ? g.n_chunks : 1, st));
    const unsigned long long init[2] = {~0ull, 0ull};
    OBJ_CU(cudaMemcpyAsync(g.status, init, 16, cudaMemcpyHostToDevice, st));
    const unsigned blocks = (nc1 + 127) / 128;
    obj_count_kernel<<<blocks, 128, 0, st>>>(g.d_text, n, g.n_chunks, counts);
    OBJ_CU(cub::DeviceScan::ExclusiveScan(g.scan_temp, g.scan_temp_bytes, counts, counts, ObjCountsSum(), zero, (int)nc1, st));
    ObjCounts total;
    OBJ_CU(cudaMemcpyAsync(&total, counts + g.n_chunks, sizeof(ObjCounts), cudaMemcpyDeviceToHost, st));
    OBJ_CU(cudaStreamSynchronize(st));
    g.n_v = total.v; g.n_vn = total.vn; g.n_vt = total.vt; g.n_f = total.f; g.n_um = total.um;
    g.launches += 2;
    for (int pass = 0; pass < 2; pass++) {
        Carver c(pass ? g.arena_stmt->base : nullptr);
        c.take(g.v, 3ull * g.n_v); c.take(g.vn, 3ull * g.n_vn); c.take(g.vt, 2ull * g.n_vt);
        c.take(g.fidx, 9ull * g.n_f); c.take(g.f_um, g.n_f); c.take(g.f_off, g.n_f);
        c.take(g.um_off, g.n_um); c.take(g.um_len, g.n_um); c.take(g.um_mat, g.n_um);
        if (pass == 0) OBJ_CU(g.arena_stmt->reserve(c.used));
    }
    if (g.n_chunks) {
        const unsigned pb = (g.n_chunks + 127) / 128;
        obj_parse_kernel<false><<<pb, 128, 0, st>>>(g.d_text, n, g.n_chunks, counts, g.defer, g.v, g.vn, g.vt, g.fidx, g.f_um, g.f_off, g.um_off, g.um_len, g.status);
        g.launches++;
        OBJ_CU(cudaMemcpyAsync(g.h_status, g.status, 16, cudaMemcpyDeviceToHost, st));
        OBJ_CU(cudaStreamSynchronize(st));
        if (g.h_status[1]) {                            // some chunk holds a number for the 896-bit conversion
            obj_parse_kernel<true><<<pb, 128, 0, st>>>(g.d_text, n, g.n_chunks, counts, g.defer, g.v, g.vn, g.vt, g.fidx, g.f_um, g.f_off, g.um_off, g.um_len, g.status);
            g.launches++;
            OBJ_CU(cudaMemcpyAsync(g.h_status, g.status, 16, cudaMemcpyDeviceToHost, st));
            OBJ_CU(cudaStreamSynchronize(st));
        }
    }
    return (int)cudaGetLastError();
}

int obj_ingest_names(ObjIngest& g, std::vector<std::string>& names, cudaStream_t st) {
    names.clear();
    if (!g.n_um) return 0;
    std::vector<unsigned> len(g.n_um), dst(g.n_um);
    OBJ_CU(cudaMemcpyAsync(len.data(), g.um_len, 4ull * g.n_um, cudaMemcpyDeviceToHost, st));
    OBJ_CU(cudaStreamSynchronize(st));
    size_t total = 0;
    for (unsigned i = 0; i < g.n_um; i++) { dst[i] = (unsigned)total; total += len[i]; }
    unsigned* d_dst = nullptr; char* d_out = nullptr;
    OBJ_CU(cudaMalloc((void**)&d_dst, 4ull * g.n_um)); OBJ_CU(cudaMalloc((void**)&d_out, total ? total : 1));
    OBJ_CU(cudaMemcpyAsync(d_dst, dst.data(), 4ull * g.n_um, cudaMemcpyHostToDevice, st));
    obj_pack_names_kernel<<<(g.n_um + 127) / 128, 128, 0, st>>>(g.d_text, g.n_um, g.um_off, g.um_len, d_dst, d_out);
    g.launches++;
    std::vector<char> packed(total ? total : 1);
    OBJ_CU(cudaMemcpyAsync(packed.data(), d_out, total, cudaMemcpyDeviceToHost, st));
    OBJ_CU(cudaStreamSynchronize(st));
    cudaFree(d_dst); cudaFree(d_out);
    names.reserve(g.n_um);
    for (unsigned i = 0; i < g.n_um; i++) names.emplace_back(packed.data() + dst[i], len[i]);
    return 0;
}

int obj_ingest_flatten(ObjIngest& g, const int* um_mat, double* d_pos, double* d_nrm, double* d_uv, int* d_mat_id, cudaStream_t st) {
    if (g.n_um) OBJ_CU(cudaMemcpyAsync(g.um_mat, um_mat, 4ull * g.n_um, cudaMemcpyHostToDevice, st));
    const unsigned long long threads = 3ull * g.n_f;
    obj_flatten_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(g.n_f, g.fidx, g.f_um, g.f_off, g.um_mat, g.v, g.vn, g.vt, d_pos, d_nrm, d_uv, d_mat_id, g.status);
    g.launches++;
    OBJ_CU(cudaMemcpyAsync(g.h_status, g.status, 16, cudaMemcpyDeviceToHost, st));
    OBJ_CU(cudaStreamSynchronize(st));
    return (int)cudaGetLastError();
}

int launch_gather_faces(const int* d_idx, int m, const double* d_pos, const double* d_nrm, double* d_out, cudaStream_t st) {
    if (m <= 0) return 0;
    gather_faces_kernel<<<(m * 18 + 255) / 256, 256, 0, st>>>(d_idx, m, d_pos, d_nrm, d_out);
    return (int)cudaGetLastError();
}

}  // namespace mcpt
