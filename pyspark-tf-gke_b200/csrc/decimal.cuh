// Decimal text -> binary64, correctly rounded (round-half-even), as Java's
// Double.parseDouble and Python's float() are - the conversion behind
// spark.read.csv(inferSchema=True) (spark_workload_to_cloud_k8s.py:48) and
// float(v) in load_csv (train_tf_ps.py:132).
//
// Exact paths only; anything else is reported, never approximated:
//   A/B  <= 2^53 mantissa, |exp10| <= 22: one IEEE multiply / divide of two
//        exactly representable numbers (Clinger's fast path);
//   C    any <= 19-digit mantissa, |exp10| <= 19: exact 128-bit integer
//        product / quotient with remainder, rounded half-even by hand.
// Outside that domain (more than 19 significant digits with a non-zero tail, or
// |exp10| > 19 beyond path A/B) the result is DEC_INEXACT and the caller fails
// loudly.  Shared by the CUDA kernels and, compiled as plain C++, by the CPU
// unit test that checks it against strtod on millions of numerals.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define DEC_HD __host__ __device__ __forceinline__
#else
#define DEC_HD inline
#endif

namespace etl {

enum DecStatus : int {
    DEC_OK = 0,       // *out holds the value (may be +-inf or NaN from "inf"/"nan" text)
    DEC_EMPTY = 1,    // nothing but whitespace
    DEC_INVALID = 2,  // not a numeral (Python float() would raise ValueError)
    DEC_INEXACT = 3,  // a numeral, but outside the exactly-converted domain
};
// bit flags returned through *flags
constexpr int DEC_FLAG_NONINT = 1;  // had '.', an exponent, or was inf/nan: column is not an integer column
constexpr int DEC_FLAG_NAN = 2;     // the text was a NaN literal

typedef unsigned __int128 dec_u128;

DEC_HD double dec_pow10(int e) {  // exact for 0 <= e <= 22: every power up to 1e22 is representable,
    double r = 1.0;               // so each product below is exact (no table: no local-memory array on the GPU)
    for (int i = 0; i < e; ++i) r *= 10.0;
    return r;
}
DEC_HD uint64_t dec_pow10_u64(int e) {  // 0 <= e <= 19
    uint64_t r = 1;
    for (int i = 0; i < e; ++i) r *= 10u;
    return r;
}
DEC_HD int dec_clz64(uint64_t x) {
#ifdef __CUDA_ARCH__
    return __clzll((long long)x);
#else
    return __builtin_clzll(x);
#endif
}
DEC_HD int dec_bits128(dec_u128 x) {
    const uint64_t hi = (uint64_t)(x >> 64), lo = (uint64_t)x;
    return hi ? 128 - dec_clz64(hi) : (lo ? 64 - dec_clz64(lo) : 0);
}
// q * 2^e2 for q <= 2^53 (exact scaling; the domain keeps it far from overflow / subnormals)
DEC_HD double dec_ldexp(uint64_t q, int e2) {
    double r = (double)q;
    while (e2 > 60) { r *= 1152921504606846976.0; e2 -= 60; }
    while (e2 < -60) { r *= (1.0 / 1152921504606846976.0); e2 += 60; }
    if (e2 >= 0) return r * (double)((uint64_t)1 << e2);
    return r * (1.0 / (double)((uint64_t)1 << -e2));
}
// round the exact value (Q + sticky) * 2^e2 to binary64, half-even
DEC_HD double dec_round128(dec_u128 Q, bool sticky, int e2) {
    const int nbits = dec_bits128(Q);
    if (nbits <= 53) return dec_ldexp((uint64_t)Q, e2);  // sticky implies nbits is large, see callers
    const int shift = nbits - 53;
    uint64_t q = (uint64_t)(Q >> shift);
    const dec_u128 rem = Q & ((((dec_u128)1) << shift) - 1);
    const dec_u128 half = ((dec_u128)1) << (shift - 1);
    if (rem > half || (rem == half && (sticky || (q & 1)))) ++q;
    return dec_ldexp(q, e2 + shift);
}

DEC_HD bool dec_is_space(uint8_t c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n' || c == '\v' || c == '\f'; }
DEC_HD uint8_t dec_lower(uint8_t c) { return (c >= 'A' && c <= 'Z') ? (uint8_t)(c + 32) : c; }

DEC_HD bool dec_match_ci(const uint8_t* p, int len, const char* word, int wlen) {
    if (len != wlen) return false;
    for (int i = 0; i < wlen; ++i)
        if (dec_lower(p[i]) != (uint8_t)word[i]) return false;
    return true;
}

// Parse p[0..len).  Leading/trailing ASCII whitespace is ignored (load_csv strips
// cells, train_tf_ps.py:128; Java's parseDouble trims too).
DEC_HD int dec_parse(const uint8_t* p, int len, double* out, int* flags) {
    *flags = 0;
    while (len > 0 && dec_is_space(p[0])) { ++p; --len; }
    while (len > 0 && dec_is_space(p[len - 1])) --len;
    if (len == 0) return DEC_EMPTY;
    int i = 0;
    bool neg = false;
    if (p[0] == '+' || p[0] == '-') { neg = p[0] == '-'; i = 1; }
    if (i == len) return DEC_INVALID;
    // inf / infinity / nan, any case (Python float(); Java accepts "NaN", "Infinity")
    if (dec_lower(p[i]) == 'i' || dec_lower(p[i]) == 'n') {
        *flags = DEC_FLAG_NONINT;
        if (dec_match_ci(p + i, len - i, "inf", 3) || dec_match_ci(p + i, len - i, "infinity", 8)) {
            const uint64_t bits = neg ? 0xfff0000000000000ull : 0x7ff0000000000000ull;
            *out = *reinterpret_cast<const double*>(&bits);
            return DEC_OK;
        }
        if (dec_match_ci(p + i, len - i, "nan", 3)) {
            const uint64_t bits = neg ? 0xfff8000000000000ull : 0x7ff8000000000000ull;
            *out = *reinterpret_cast<const double*>(&bits);
            *flags |= DEC_FLAG_NAN;
            return DEC_OK;
        }
        return DEC_INVALID;
    }
    uint64_t m = 0;       // first <= 19 significant digits
    int sig = 0;          // significant digits taken into m
    int dropped = 0;      // integer-part digits not taken (scale up by 10^dropped)
    bool tail_nonzero = false;
    int frac_taken = 0;   // fraction digits taken into m
    int ndigits = 0;
    for (; i < len && p[i] >= '0' && p[i] <= '9'; ++i) {
        ++ndigits;
        const unsigned d = p[i] - '0';
        if (sig == 0 && d == 0) continue;  // leading zeros
        if (sig < 19) { m = m * 10 + d; ++sig; }
        else { ++dropped; tail_nonzero |= d != 0; }
    }
    if (i < len && p[i] == '.') {
        ++i;
        *flags |= DEC_FLAG_NONINT;
        for (; i < len && p[i] >= '0' && p[i] <= '9'; ++i) {
            ++ndigits;
            const unsigned d = p[i] - '0';
            if (sig == 0 && d == 0) { ++frac_taken; continue; }  // 0.000x: scales, not significant
            if (sig < 19) { m = m * 10 + d; ++sig; ++frac_taken; }
            else tail_nonzero |= d != 0;
        }
    }
    if (ndigits == 0) return DEC_INVALID;
    long long e10 = (long long)dropped - frac_taken;
    if (i < len && (p[i] == 'e' || p[i] == 'E')) {
        ++i;
        *flags |= DEC_FLAG_NONINT;
        bool eneg = false;
        if (i < len && (p[i] == '+' || p[i] == '-')) { eneg = p[i] == '-'; ++i; }
        if (i == len) return DEC_INVALID;
        long long ev = 0;
        int ed = 0;
        for (; i < len && p[i] >= '0' && p[i] <= '9'; ++i) {
            if (ev < 100000) ev = ev * 10 + (p[i] - '0');
            ++ed;
        }
        if (ed == 0) return DEC_INVALID;
        e10 += eneg ? -ev : ev;
    }
    if (i != len) return DEC_INVALID;
    double r;
    if (m == 0) {
        r = 0.0;
    } else if (tail_nonzero) {
        return DEC_INEXACT;
    } else if (m <= ((uint64_t)1 << 53) && e10 >= -22 && e10 <= 22) {
        r = e10 >= 0 ? (double)m * dec_pow10((int)e10) : (double)m / dec_pow10((int)-e10);
    } else if (e10 >= 0 && e10 <= 19) {
        r = dec_round128((dec_u128)m * dec_pow10_u64((int)e10), false, 0);
    } else if (e10 < 0 && e10 >= -19) {
        const int lz = dec_clz64(m);
        const dec_u128 N = ((dec_u128)(m << lz)) << 64;
        const uint64_t d = dec_pow10_u64((int)-e10);
        r = dec_round128(N / d, (N % d) != 0, -(64 + lz));
    } else {
        return DEC_INEXACT;
    }
    *out = neg ? -r : r;
    return DEC_OK;
}

}  // namespace etl
