// png_device.cu — the output stage on the GPU: imshow's quantisation
// (uchar)clamp(v*255, 0, 255) (MTPC/MTPC.cpp:22-30) and the uncompressed-PNG framing that MTPC/svpng.inc:77-108 emits
// with one fputc per byte — the same bytes svpng writes (one stored block per scanline; tests compare with a file the
// reference wrote) and the host writer (host/png_writer.cpp).  At 3840x2160 the host stage
// costs as much as an 8-GPU render of the image; here the float image never leaves the device: the D2H copy is the
// finished PNG (1 byte per channel instead of 4).
//   K13 png_pack_kernel   one thread per raw scanline byte: filter byte / quantised channel into its place inside the
//                         zlib stored blocks, block headers, and the two integer sums Adler-32 is made of
//   K14 png_crc_kernel    CRC-32 of the IDAT chunk: every thread runs the table-driven CRC over its own 2 KiB segment,
//                         shifts it past the bytes that follow (multiplication by x^(8n) mod P, zlib's crc32_combine
//                         identity) and XORs it into the total
//   K15 png_finish_kernel one thread: signature, IHDR, IDAT length/tag, zlib header, Adler-32, CRCs, IEND
#include <cstdint>
#include "render.h"

namespace mcpt {

#define PNG_SEG 2048u
#define PNG_POLY 0xedb88320u

struct PngLayout {
    unsigned long long raw;      // (3w+1)*h scanline bytes
    unsigned long long nblocks;  // stored deflate blocks: ONE PER SCANLINE, as svpng writes them (MTPC/svpng.inc:98-105)
    unsigned long long zsize;    // 2 + h*(5 + 3w+1) + 4  (svpng.inc:96)
    unsigned long long total;    // 8 + 25 + 12 + zsize + 12
    unsigned long long idat_tag; // offset of "IDAT"
    unsigned long long zoff;     // offset of the zlib stream
};

__host__ __device__ inline PngLayout png_layout(int w, int h) {
    PngLayout L;
    L.raw = (3ull * (unsigned long long)w + 1ull) * (unsigned long long)h;
    L.nblocks = (unsigned long long)h;
    L.zsize = 2ull + 5ull * L.nblocks + L.raw + 4ull;
    L.idat_tag = 8ull + 25ull + 4ull;
    L.zoff = L.idat_tag + 4ull;
    L.total = L.zoff + L.zsize + 4ull + 12ull;
    return L;
}

// sums[0] = sum of raw bytes, sums[1] = sum of (raw - idx) * byte  (Adler-32: A = 1 + S1, B = raw + S2, both mod 65521)
__global__ void __launch_bounds__(256) png_pack_kernel(const float* __restrict__ rgb, int w, int h, unsigned char* __restrict__ png,
                                                       unsigned long long* __restrict__ sums) {
    const PngLayout L = png_layout(w, h);
    const unsigned long long row = 3ull * w + 1ull;
    unsigned long long s1 = 0, s2 = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < L.raw; i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long y = i / row, c = i - y * row;
        unsigned v = 0;
        if (c != 0) {
            double d = (double)rgb[y * (row - 1ull) + (c - 1ull)] * 255.0;    // the reference scales a double (image::img)
            d = d < 0.0 ? 0.0 : (d > 255.0 ? 255.0 : d);
            v = d == d ? (unsigned)d : 0u;                                    // truncation; NaN -> 0
        }
        unsigned char* z = png + L.zoff + 2ull + 5ull * (y + 1ull) + i;   // scanline y sits behind y+1 block headers
        *z = (unsigned char)v;
        if (c == 0) {   // the filter byte's thread also writes its scanline's 5-byte stored-block header
            z[-5] = (y + 1ull == (unsigned long long)h) ? 1 : 0;
            z[-4] = (unsigned char)row; z[-3] = (unsigned char)(row >> 8);
            z[-2] = (unsigned char)(~row); z[-1] = (unsigned char)((~row) >> 8);
        }
        // the weight is reduced mod 65521 first, so the 64-bit sums cannot overflow for any image the 32-bit IDAT length allows
        s1 += v; s2 += ((L.raw - i) % 65521ull) * v;
    }
    // integer sums: order-independent, so the block reduction + atomics are deterministic
    for (int off = 16; off > 0; off >>= 1) { s1 += __shfl_down_sync(0xffffffffu, s1, off); s2 += __shfl_down_sync(0xffffffffu, s2, off); }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&sums[0], s1); atomicAdd(&sums[1], s2); }
}

// a * b mod P in the reflected representation (zlib crc32.c multmodp)
__device__ __forceinline__ unsigned multmodp(unsigned a, unsigned b) {
    unsigned m = 1u << 31, p = 0;
    for (;;) {
        if (a & m) { p ^= b; if ((a & (m - 1u)) == 0u) break; }
        m >>= 1;
        b = (b & 1u) ? (b >> 1) ^ PNG_POLY : b >> 1;
    }
    return p;
}
// x^(n * 2^k) mod P
__device__ __forceinline__ unsigned x2nmodp(unsigned long long n, unsigned k, const unsigned* x2n) {
    unsigned p = 1u << 31;
    while (n) {
        if (n & 1ull) p = multmodp(x2n[k & 31u], p);
        n >>= 1; k++;
    }
    return p;
}

// crc_out ^= shift(crc32(segment), bytes after the segment) over [begin, begin + len)
__global__ void __launch_bounds__(128) png_crc_kernel(const unsigned char* __restrict__ png, unsigned long long begin,
                                                      unsigned long long len, unsigned* __restrict__ crc_out) {
    __shared__ unsigned table[256];
    __shared__ unsigned x2n[32];
    for (unsigned n = threadIdx.x; n < 256; n += blockDim.x) {
        unsigned c = n;
        for (int k = 0; k < 8; k++) c = (c & 1u) ? PNG_POLY ^ (c >> 1) : c >> 1;
        table[n] = c;
    }
    if (threadIdx.x == 0) {
        unsigned p = 0x40000000u;   // x^1
        x2n[0] = p;
        for (int n = 1; n < 32; n++) { p = multmodp(p, p); x2n[n] = p; }
    }
    __syncthreads();
    const unsigned long long nseg = (len + PNG_SEG - 1) / PNG_SEG;
    unsigned acc = 0;
    for (unsigned long long sgi = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; sgi < nseg; sgi += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long off = sgi * PNG_SEG, n = len - off < PNG_SEG ? len - off : PNG_SEG;
        const unsigned char* p = png + begin + off;
        unsigned c = 0xFFFFFFFFu;
        for (unsigned long long i = 0; i < n; i++) c = table[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
        c = ~c;
        const unsigned long long after = len - off - n;
        acc ^= after ? multmodp(x2nmodp(after, 3u, x2n), c) : c;
    }
    for (int off = 16; off > 0; off >>= 1) acc ^= __shfl_down_sync(0xffffffffu, acc, off);
    if ((threadIdx.x & 31) == 0 && acc) atomicXor(crc_out, acc);
}

__device__ __forceinline__ void put_be32(unsigned char* p, unsigned v) { p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v; }
__device__ unsigned crc_small(const unsigned char* p, int n) {
    unsigned c = 0xFFFFFFFFu;
    for (int i = 0; i < n; i++) { c ^= p[i]; for (int k = 0; k < 8; k++) c = (c & 1u) ? PNG_POLY ^ (c >> 1) : c >> 1; }
    return ~c;
}

// phase 0 (before the IDAT CRC): everything except the IDAT CRC; phase 1: the IDAT CRC
__global__ void png_finish_kernel(unsigned char* png, int w, int h, const unsigned long long* sums, const unsigned* crc, int phase) {
    const PngLayout L = png_layout(w, h);
    if (phase == 1) { put_be32(png + L.zoff + L.zsize, *crc); return; }
    const unsigned char sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
    for (int i = 0; i < 8; i++) png[i] = sig[i];
    unsigned char* ih = png + 8;
    put_be32(ih, 13u); ih[4] = 'I'; ih[5] = 'H'; ih[6] = 'D'; ih[7] = 'R';
    put_be32(ih + 8, (unsigned)w); put_be32(ih + 12, (unsigned)h);
    ih[16] = 8; ih[17] = 2; ih[18] = 0; ih[19] = 0; ih[20] = 0;
    put_be32(ih + 21, crc_small(ih + 4, 17));
    unsigned char* id = png + 8 + 25;
    put_be32(id, (unsigned)L.zsize); id[4] = 'I'; id[5] = 'D'; id[6] = 'A'; id[7] = 'T';
    png[L.zoff] = 0x78; png[L.zoff + 1] = 0x01;
    const unsigned a = (unsigned)((1ull + sums[0]) % 65521ull), b = (unsigned)((L.raw + sums[1]) % 65521ull);
    put_be32(png + L.zoff + L.zsize - 4, (b << 16) | a);
    unsigned char* ie = png + L.zoff + L.zsize + 4;
    put_be32(ie, 0u); ie[4] = 'I'; ie[5] = 'E'; ie[6] = 'N'; ie[7] = 'D';
    put_be32(ie + 8, crc_small(ie + 4, 4));
}

size_t png_encoded_size(int w, int h) { return (size_t)png_layout(w, h).total; }

// d_png: png_encoded_size bytes; d_scratch: 32 bytes (two 64-bit sums + the CRC word).  4 launches on `stream`.
int launch_png_encode(const float* d_rgb, int w, int h, unsigned char* d_png, void* d_scratch, cudaStream_t stream) {
    const PngLayout L = png_layout(w, h);
    if (L.zsize > 0xffffffffull || 3ull * w + 1ull > 65535ull) return (int)cudaErrorInvalidValue;   // IDAT length is 32 bits, a block's LEN 16
    cudaError_t e = cudaMemsetAsync(d_scratch, 0, 32, stream);
    if (e != cudaSuccess) return (int)e;
    unsigned long long* sums = (unsigned long long*)d_scratch;
    unsigned* crc = (unsigned*)((char*)d_scratch + 16);
    const unsigned long long want = (L.raw + 255ull) / 256ull;
    const int blocks = (int)(want < 148ull * 16ull ? want : 148ull * 16ull);
    png_pack_kernel<<<blocks, 256, 0, stream>>>(d_rgb, w, h, d_png, sums);
    png_finish_kernel<<<1, 1, 0, stream>>>(d_png, w, h, sums, crc, 0);
    const unsigned long long clen = 4ull + L.zsize, nseg = (clen + PNG_SEG - 1) / PNG_SEG;
    const unsigned long long cwant = (nseg + 127ull) / 128ull;
    png_crc_kernel<<<(int)(cwant < 148ull * 8ull ? cwant : 148ull * 8ull), 128, 0, stream>>>(d_png, L.idat_tag, clen, crc);
    png_finish_kernel<<<1, 1, 0, stream>>>(d_png, w, h, sums, crc, 1);
    return (int)cudaGetLastError();
}

// ---- mcpt_encode_png_static: the output stage for callers that must not allocate ------------------------------------------
// No counterpart in the reference.  mcpt_encode_png keeps its device-side PNG
// buffer in a context and grows it with cudaMalloc; code under CUDA graph capture, tight serving loops and harnesses that forbid
// allocation cannot call that.  This variant keeps the same scratch — 32 B of checksum accumulators (which must start at zero:
// cleared by a cudaMemsetAsync on the caller's stream every call) and the encoded file before it is copied out — in ONE
// file-scope device array of MCPT_PNG_STATIC_BYTES (8 MiB: up to 1920 x 1080; every GPU that loads libmcpt.so holds it until
// exit).  It needs no mcpt_ctx and calls no cudaMalloc / cudaFree / allocating library routine.  Every call on a device shares
// the array, which keeps the last call's contents: calls that overlap on the GPU (different streams, different threads) RACE.
// In host code `g_png_static` is a host-side stand-in, not a device address: the address is asked of the runtime once per device,
// with that device current (cudaGetSymbolAddress), and that pointer is what the kernels get.
#define PNG_STATIC_HEADER 256   /* the accumulators; keeps the PNG bytes 256-B aligned */
static __device__ __align__(256) unsigned char g_png_static[MCPT_PNG_STATIC_BYTES];

}  // namespace mcpt (closed for the C entry point below)

#include <mutex>
#include <string>
void mcpt_set_error(const char* msg);

extern "C" MCPT_API int mcpt_encode_png_static(int device, const float* d_rgb, int width, int height, uint8_t* png_out, int64_t capacity,
                                               void* cuda_stream) {
    using namespace mcpt;
    auto fail = [](int code, const std::string& m) { mcpt_set_error(m.c_str()); return code; };
    if (!d_rgb || !png_out || width <= 0 || height <= 0) return fail(MCPT_ERR_ARG, "mcpt_encode_png_static: bad argument");
    if (width > 21844) return fail(MCPT_ERR_ARG, "mcpt_encode_png_static: width > 21844 does not fit svpng's one-block-per-scanline framing");
    const unsigned long long bytes = png_layout(width, height).total;
    if (bytes + PNG_STATIC_HEADER > (unsigned long long)MCPT_PNG_STATIC_BYTES)
        return fail(MCPT_ERR_ARG, "mcpt_encode_png_static: a " + std::to_string(width) + " x " + std::to_string(height) + " image encodes to " +
                                      std::to_string(bytes) + " bytes, more than the static scratch holds (" +
                                      std::to_string((unsigned long long)MCPT_PNG_STATIC_BYTES - PNG_STATIC_HEADER) +
                                      " bytes, e.g. 1920 x 1080); use mcpt_encode_png");
    if (capacity < (int64_t)bytes) return fail(MCPT_ERR_ARG, "mcpt_encode_png_static: output buffer smaller than mcpt_png_size()");
    enum { MAX_DEV = 64 };
    if (device < 0 || device >= MAX_DEV) return fail(MCPT_ERR_ARG, "mcpt_encode_png_static: device index out of range");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(MCPT_ERR_CUDA, std::string("mcpt_encode_png_static: cudaSetDevice: ") + cudaGetErrorName(e));
    // set-up, once per device, with that device current: the DEVICE address of the array
    static std::mutex mu;
    static unsigned char* dev_ptr[MAX_DEV] = {};
    unsigned char* base = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!dev_ptr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_png_static);
            if (e != cudaSuccess || !p) return fail(MCPT_ERR_CUDA, std::string("mcpt_encode_png_static: cudaGetSymbolAddress: ") + cudaGetErrorName(e));
            dev_ptr[device] = (unsigned char*)p;
        }
        base = dev_ptr[device];
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int le = launch_png_encode(d_rgb, width, height, base + PNG_STATIC_HEADER, base, st);   // clears the accumulators first
    if (le) return fail(MCPT_ERR_CUDA, std::string("mcpt_encode_png_static: launch: ") + cudaGetErrorName((cudaError_t)le));
    e = cudaMemcpyAsync(png_out, base + PNG_STATIC_HEADER, (size_t)bytes, cudaMemcpyDefault, st);   // png_out: host or device
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return fail(MCPT_ERR_CUDA, std::string("mcpt_encode_png_static: copy out: ") + cudaGetErrorName(e));
    return MCPT_OK;
}
