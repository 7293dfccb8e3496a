// Philox4x32-10 (Salmon et al., SC'11) counter-based RNG.  Keyed by (pixel, sample) with counter
// (bounce, stream, seed_lo, seed_hi): a sample's random numbers depend on nothing else, so images
// are independent of launch shape and GPU count.  oracle/mcpt_oracle.c uses the same mapping.
#pragma once
#include <cuda_runtime.h>

namespace mcpt {

enum { STREAM_PATH = 0, STREAM_FRESNEL = 1, STREAM_LIGHT0 = 2, STREAM_JITTER = 0x7FFFFFFF };   // LIGHT0 + l for light l

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const unsigned hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const unsigned hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}
// 24-bit uniforms in [0,1): exactly representable in fp32, identical to the oracle's doubles.
__device__ __forceinline__ float4 philox_uniform4(unsigned bounce, unsigned stream, unsigned seed_lo, unsigned seed_hi,
                                                  unsigned pixel, unsigned sample) {
    const uint4 r = philox4x32_10(make_uint4(bounce, stream, seed_lo, seed_hi), make_uint2(pixel, sample));
    const float s = 1.0f / 16777216.0f;
    return make_float4((r.x >> 8) * s, (r.y >> 8) * s, (r.z >> 8) * s, (r.w >> 8) * s);
}

}  // namespace mcpt
