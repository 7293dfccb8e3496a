// obj_number.cuh — correctly rounded decimal -> double for the GPU scene ingest (obj_ingest.cu).
//
// The reference reads every OBJ number with atof (MTPC/sceneManagement.cpp:87-92,103-108,119-123), i.e. glibc strtod:
// the longest valid decimal prefix of the token, rounded to nearest-even.  The ingest must reproduce those bits, so
// this is an exact conversion, not an approximation:
//   fast    <= 15-16 digits (w <= 2^53) and |e10| <= 22: one IEEE multiply or divide by an exact power of ten (Clinger)
//   medium  <= 19 digits and |e10| <= 27: w * 5^e exactly in 128 bits, or a 65-step restoring division w / 5^-e, then one
//           round-half-even of the 64-bit quotient + sticky bit
//   big     <= 40 digits, any exponent: the same algorithm on 896-bit integers (rare; run by a second launch over the
//           chunks that asked for it, so its local-memory arrays never weigh on the main kernel)
// Tokens glibc would read differently are REPORTED, never guessed: hex floats, inf/nan, more than 40 significant digits.
// Plain C++ apart from the __host__ __device__ markers: tests/test_obj_number.py compiles it with g++ and checks it
// against strtod on the host.
#pragma once
#include <cstdint>

#ifdef __CUDACC__
#define MCPT_HD __host__ __device__
#else
#define MCPT_HD
#endif

namespace mcpt {
namespace objnum {

typedef unsigned long long u64;
typedef unsigned __int128 u128;

enum { NUM_OK = 0, NUM_DEFER = 1, NUM_UNSUPPORTED = 2 };
#define MCPT_BIG_LIMBS 14
#define MCPT_BIG_DIGITS 40

MCPT_HD inline int clz64(u64 v) {
#ifdef __CUDA_ARCH__
    return __clzll((long long)v);
#else
    return v ? __builtin_clzll(v) : 64;
#endif
}
MCPT_HD inline double bits_to_double(u64 b) {
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)b);
#else
    union { u64 u; double d; } x; x.u = b; return x.d;
#endif
}

// 5^k, k <= 27 (5^27 < 2^63); built by repeated multiplication so that no table has to live in constant memory
MCPT_HD inline u64 pow5_u64(int k) {
    u64 p = 1, b = 5;
    while (k) { if (k & 1) p *= b; b *= b; k >>= 1; }
    return p;
}
MCPT_HD inline double pow10_exact(int k) {   // k <= 22: 10^k is a double
    double p = 1.0, b = 10.0;
    while (k) { if (k & 1) p *= b; b *= b; k >>= 1; }   // every partial product is <= 10^22 < 2^53 * 2^k: exact
    return p;
}

// value = q * 2^x with bit 63 of q set (+ an infinitesimal if sticky): round to nearest-even double
MCPT_HD inline double round_pack(u64 q, bool sticky, int x, bool neg) {
    const u64 sign = neg ? 0x8000000000000000ull : 0ull;
    int E = x + 63;                                   // unbiased exponent of the leading bit
    if (E > 1023) return bits_to_double(sign | 0x7FF0000000000000ull);
    int shift = 11;
    if (E < -1022) shift += -1022 - E;                // subnormal: fewer mantissa bits
    u64 mant, rem, half;
    if (shift > 64) return bits_to_double(sign);      // below half of the smallest subnormal
    if (shift == 64) { mant = 0; rem = q; half = 0x8000000000000000ull; }
    else { mant = q >> shift; rem = q & ((1ull << shift) - 1ull); half = 1ull << (shift - 1); }
    if (rem > half || (rem == half && (sticky || (mant & 1ull)))) mant++;
    if (E < -1022) return bits_to_double(sign | mant);   // mant == 2^52 is the smallest normal: the encoding carries
    if (mant >> 53) { mant >>= 1; E++; if (E > 1023) return bits_to_double(sign | 0x7FF0000000000000ull); }
    return bits_to_double(sign | ((u64)(E + 1023) << 52) | (mant & 0x000FFFFFFFFFFFFFull));
}

// w * 10^e10, w != 0, |e10| <= 27, exactly rounded
MCPT_HD inline double scale_u64(u64 w, int e10, bool neg) {
    u64 q; bool sticky; int x;
    if (e10 >= 0) {
        const u128 N = (u128)w * pow5_u64(e10);
        const u64 hi = (u64)(N >> 64), lo = (u64)N;
        const int bl = hi ? 128 - clz64(hi) : 64 - clz64(lo);
        if (bl <= 64) { q = lo << (64 - bl); sticky = false; }
        else { const int sh = bl - 64; q = (u64)(N >> sh); sticky = (N & (((u128)1 << sh) - 1)) != 0; }
        x = bl - 64 + e10;
    } else {
        const int m = -e10;
        const u64 P = pow5_u64(m);
        const int lzw = clz64(w), lzp = clz64(P);
        u128 R = (u128)(w << lzw);
        const u128 D = (u128)(P << lzp);
        int t = 0;
        q = 0;
        do {                                          // R < 2D throughout: one quotient bit per step
            const bool bit = R >= D;
            if (bit) R -= D;
            q = (q << 1) | (bit ? 1ull : 0ull);
            R <<= 1;
            t++;
        } while (!(q >> 63));
        sticky = R != 0;
        x = -(t - 1) + lzp - lzw - m;
    }
    return round_pack(q, sticky, x, neg);
}

// ---- 896-bit path -------------------------------------------------------------------------------------------------
struct Big { u64 d[MCPT_BIG_LIMBS]; };
MCPT_HD inline void big_set(Big& a, u64 v) { a.d[0] = v; for (int i = 1; i < MCPT_BIG_LIMBS; i++) a.d[i] = 0; }
MCPT_HD inline void big_mul_add(Big& a, u64 m, u64 add) {
    u64 carry = add;
    for (int i = 0; i < MCPT_BIG_LIMBS; i++) { const u128 t = (u128)a.d[i] * m + carry; a.d[i] = (u64)t; carry = (u64)(t >> 64); }
}
MCPT_HD inline int big_bitlen(const Big& a) {
    for (int i = MCPT_BIG_LIMBS - 1; i >= 0; i--) if (a.d[i]) return 64 * i + 64 - clz64(a.d[i]);
    return 0;
}
MCPT_HD inline void big_shl(Big& a, int k) {
    const int ls = k >> 6, bs = k & 63;
    for (int i = MCPT_BIG_LIMBS - 1; i >= 0; i--) {
        u64 v = 0;
        if (i - ls >= 0) v = a.d[i - ls] << bs;
        if (bs && i - ls - 1 >= 0) v |= a.d[i - ls - 1] >> (64 - bs);
        a.d[i] = v;
    }
}
MCPT_HD inline bool big_ge(const Big& a, const Big& b) {
    for (int i = MCPT_BIG_LIMBS - 1; i >= 0; i--) if (a.d[i] != b.d[i]) return a.d[i] > b.d[i];
    return true;
}
MCPT_HD inline void big_sub(Big& a, const Big& b) {
    u64 borrow = 0;
    for (int i = 0; i < MCPT_BIG_LIMBS; i++) {
        const u64 x = a.d[i], y = b.d[i];
        const u64 r = x - y - borrow;
        borrow = (x < y || (x == y && borrow)) ? 1ull : 0ull;
        a.d[i] = r;
    }
}
MCPT_HD inline bool big_is_zero(const Big& a) { for (int i = 0; i < MCPT_BIG_LIMBS; i++) if (a.d[i]) return false; return true; }
MCPT_HD inline void big_pow5(Big& p, int m) {
    big_set(p, 1);
    while (m >= 27) { big_mul_add(p, 7450580596923828125ull, 0); m -= 27; }   // 5^27
    if (m) big_mul_add(p, pow5_u64(m), 0);
}

// N * 10^e10 (N != 0, N < 10^40), exactly rounded
MCPT_HD inline double scale_big(Big& N, int e10, bool neg) {
    if (e10 > 310) return round_pack(0x8000000000000000ull, false, 2000, neg);   // inf
    if (e10 < -370) return round_pack(0x8000000000000000ull, false, -2000, neg); // 0
    u64 q; bool sticky; int x;
    if (e10 >= 0) {
        int m = e10;
        while (m >= 27) { big_mul_add(N, 7450580596923828125ull, 0); m -= 27; }
        if (m) big_mul_add(N, pow5_u64(m), 0);
        const int bl = big_bitlen(N);
        if (bl <= 64) { q = N.d[0] << (64 - bl); sticky = false; }
        else {
            const int sh = bl - 64, limb = sh >> 6, bs = sh & 63;
            q = N.d[limb] >> bs;
            if (bs) q |= N.d[limb + 1] << (64 - bs);
            sticky = bs && (N.d[limb] & ((1ull << bs) - 1ull));
            for (int i = 0; i < limb; i++) sticky = sticky || N.d[i];
        }
        x = bl - 64 + e10;
    } else {
        const int m = -e10;
        Big D;
        big_pow5(D, m);
        const int bn = big_bitlen(N), bp = big_bitlen(D);
        const int sh = bn - bp;                        // N / P = (R / D) * 2^sh with bitlen(R) == bitlen(D)
        if (sh >= 0) big_shl(D, sh); else big_shl(N, -sh);
        int t = 0;
        q = 0;
        do {
            const bool bit = big_ge(N, D);
            if (bit) big_sub(N, D);
            q = (q << 1) | (bit ? 1ull : 0ull);
            big_shl(N, 1);
            t++;
        } while (!(q >> 63));
        sticky = !big_is_zero(N);
        x = -(t - 1) + sh - m;
    }
    return round_pack(q, sticky, x, neg);
}

MCPT_HD inline bool is_digit(char c) { return c >= '0' && c <= '9'; }
MCPT_HD inline char lower(char c) { return (c >= 'A' && c <= 'Z') ? (char)(c + 32) : c; }

// atof of the token [p, end) (no blanks inside).  BIG = false: numbers that need the 896-bit path return NUM_DEFER
// (value untouched); BIG = true: they are converted.  NUM_UNSUPPORTED: glibc would read this token differently.
// It: a random-access iterator over chars (const char* on the host; the staged-text iterator of obj_ingest.cu on the device).
template <bool BIG, class It>
MCPT_HD inline int parse_double(It p, const It end, double& value) {
    bool neg = false;
    if (p < end && (*p == '+' || *p == '-')) { neg = *p == '-'; p++; }
    if (end - p >= 3) {
        const char a = lower(p[0]), b = lower(p[1]), c = lower(p[2]);
        if ((a == 'i' && b == 'n' && c == 'f') || (a == 'n' && b == 'a' && c == 'n')) return NUM_UNSUPPORTED;
    }
    if (end - p >= 2 && p[0] == '0' && lower(p[1]) == 'x') return NUM_UNSUPPORTED;
    const It mant = p;
    u64 w = 0;
    int nd = 0, e10 = 0;
    bool any = false, dropped = false;
    for (; p < end && is_digit(*p); p++) {
        any = true;
        const int d = *p - '0';
        if (nd == 0 && d == 0) continue;
        if (nd < 19) { w = w * 10 + (u64)d; nd++; } else { e10++; nd++; dropped = dropped || d != 0; }
    }
    if (p < end && *p == '.') {
        p++;
        for (; p < end && is_digit(*p); p++) {
            any = true;
            const int d = *p - '0';
            if (nd == 0 && d == 0) { e10--; continue; }
            if (nd < 19) { w = w * 10 + (u64)d; nd++; e10--; } else { nd++; dropped = dropped || d != 0; }
        }
    }
    if (!any) { value = 0.0; return NUM_OK; }          // no conversion: atof returns +0
    if (p < end && (*p == 'e' || *p == 'E')) {
        It r = p + 1;
        bool eneg = false;
        if (r < end && (*r == '+' || *r == '-')) { eneg = *r == '-'; r++; }
        if (r < end && is_digit(*r)) {
            int ex = 0;
            for (; r < end && is_digit(*r); r++) if (ex < 100000) ex = ex * 10 + (*r - '0');
            e10 += eneg ? -ex : ex;
            p = r;
        }
    }
    if (w == 0) { value = neg ? -0.0 : 0.0; return NUM_OK; }
    if (!dropped) {
        if (w <= (1ull << 53) && e10 >= -22 && e10 <= 22) {
            const double r = (double)w;
            const double s = e10 < 0 ? r / pow10_exact(-e10) : r * pow10_exact(e10);
            value = neg ? -s : s;
            return NUM_OK;
        }
        while (e10 > 27 && w <= 0xFFFFFFFFFFFFFFFFull / 10ull) { w *= 10ull; e10--; }
        if (e10 >= -27 && e10 <= 27) { value = scale_u64(w, e10, neg); return NUM_OK; }
    }
    if (!BIG) return NUM_DEFER;
    // 896-bit path: re-read the mantissa digits exactly
    Big N;
    big_set(N, 0);
    nd = 0; e10 = 0;
    It s = mant;
    bool frac = false;
    for (; s < end; s++) {
        if (*s == '.') { if (frac) break; frac = true; continue; }
        if (!is_digit(*s)) break;
        const int d = *s - '0';
        if (nd == 0 && d == 0) { if (frac) e10--; continue; }
        if (nd < MCPT_BIG_DIGITS) { big_mul_add(N, 10ull, (u64)d); nd++; if (frac) e10--; }
        else { if (d != 0) return NUM_UNSUPPORTED; if (!frac) e10++; }
    }
    if (s < end && (*s == 'e' || *s == 'E')) {
        It r = s + 1;
        bool eneg = false;
        if (r < end && (*r == '+' || *r == '-')) { eneg = *r == '-'; r++; }
        if (r < end && is_digit(*r)) {
            int ex = 0;
            for (; r < end && is_digit(*r); r++) if (ex < 100000) ex = ex * 10 + (*r - '0');
            e10 += eneg ? -ex : ex;
        }
    }
    value = scale_big(N, e10, neg);
    return NUM_OK;
}

// atol of [p, end): optional sign, digits; saturates far outside any valid index
template <class It>
MCPT_HD inline long long parse_long(It p, const It end) {
    bool neg = false;
    if (p < end && (*p == '+' || *p == '-')) { neg = *p == '-'; p++; }
    long long v = 0;
    for (; p < end && is_digit(*p); p++) if (v < 4000000000000000000ll / 10) v = v * 10 + (*p - '0');
    return neg ? -v : v;
}

}  // namespace objnum
}  // namespace mcpt
