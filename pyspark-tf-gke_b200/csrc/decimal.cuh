// Decimal text -> binary64, correctly rounded (round-half-even), as Java's
// Double.parseDouble and Python's float() are - the conversion behind
// spark.read.csv(inferSchema=True) (spark_workload_to_cloud_k8s.py:48) and
// float(v) in load_csv (train_tf_ps.py:132).
//
// Exact paths only; anything else is reported, never approximated:
//   A/B  <= 2^53 mantissa, |exp10| <= 22: one IEEE multiply / divide of two
//        exactly representable numbers (Clinger's fast path);
//   C    any <= 19-digit mantissa, |exp10| <= 19: exact 128-bit integer
//        product / quotient with remainder, rounded half-even by hand;
//   D    any <= 19-digit mantissa, any exponent (overflow to inf, underflow
//        through the subnormals to 0): exact multi-word integers - m * 5^e for
//        e >= 0, a 66-step shift-subtract division by 5^-e for e < 0 - rounded
//        half-even with a sticky bit.  Slow (thousands of integer ops) but rare.
//        A numeral with 20..60 significant digits takes the same path with a
//        multi-word mantissa.
// More than 60 significant digits with a non-zero tail is DEC_INEXACT and the
// caller fails loudly.  Shared by the CUDA kernels and, compiled as plain C++, by the CPU
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
constexpr int DEC_FLAG_NAN = 2;     // the text was the bare word nan (any case, no sign)

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

// ---- path D: multi-word exact scaling ------------------------------------------
constexpr int DEC_WORDS = 16;  // 1024 bits: 5^345 (801 bits) and m * 5^310 (784 bits) fit

struct DecBig {
    uint64_t w[DEC_WORDS];  // little endian
    int n;                  // words in use (w[n-1] != 0), 0 = zero
};
DEC_HD void big_set(DecBig& a, uint64_t v) {
    for (int i = 0; i < DEC_WORDS; ++i) a.w[i] = 0;
    a.w[0] = v;
    a.n = v ? 1 : 0;
}
DEC_HD void big_mul_small(DecBig& a, uint64_t f) {
    uint64_t carry = 0;
    for (int i = 0; i < a.n; ++i) {
        const dec_u128 p = (dec_u128)a.w[i] * f + carry;
        a.w[i] = (uint64_t)p;
        carry = (uint64_t)(p >> 64);
    }
    if (carry && a.n < DEC_WORDS) a.w[a.n++] = carry;
}
DEC_HD void big_pow5(DecBig& a, int k) {  // a *= 5^k
    const uint64_t p27 = 7450580596923828125ull;  // 5^27 < 2^63
    while (k >= 27) { big_mul_small(a, p27); k -= 27; }
    uint64_t f = 1;
    for (int i = 0; i < k; ++i) f *= 5;
    if (f > 1) big_mul_small(a, f);
}
DEC_HD int big_bits(const DecBig& a) { return a.n ? 64 * (a.n - 1) + 64 - dec_clz64(a.w[a.n - 1]) : 0; }
DEC_HD bool big_bit(const DecBig& a, int i) { return (a.w[i >> 6] >> (i & 63)) & 1; }
DEC_HD int big_cmp(const DecBig& a, const DecBig& b) {
    if (a.n != b.n) return a.n < b.n ? -1 : 1;
    for (int i = a.n - 1; i >= 0; --i)
        if (a.w[i] != b.w[i]) return a.w[i] < b.w[i] ? -1 : 1;
    return 0;
}
DEC_HD void big_sub(DecBig& a, const DecBig& b) {  // a -= b, a >= b
    uint64_t borrow = 0;
    for (int i = 0; i < a.n; ++i) {
        const uint64_t bi = i < b.n ? b.w[i] : 0;
        const uint64_t d = a.w[i] - bi - borrow;
        borrow = (a.w[i] < bi) || (a.w[i] == bi && borrow) ? 1 : 0;
        a.w[i] = d;
    }
    while (a.n && a.w[a.n - 1] == 0) --a.n;
}
DEC_HD void big_shl1_or(DecBig& a, bool bit) {  // a = (a << 1) | bit
    uint64_t carry = bit ? 1 : 0;
    for (int i = 0; i < a.n; ++i) {
        const uint64_t nc = a.w[i] >> 63;
        a.w[i] = (a.w[i] << 1) | carry;
        carry = nc;
    }
    if (carry && a.n < DEC_WORDS) a.w[a.n++] = carry;
    if (a.n == 0 && bit) { a.w[0] = 1; a.n = 1; }
}
// bits [lo, lo + 64) of a (lo may be negative: zero-filled)
DEC_HD uint64_t big_extract64(const DecBig& a, int lo) {
    uint64_t r = 0;
    for (int i = 0; i < 64; ++i) {
        const int b = lo + i;
        if (b >= 0 && b < 64 * a.n && big_bit(a, b)) r |= 1ull << i;
    }
    return r;
}
DEC_HD bool big_any_below(const DecBig& a, int lo) {  // any set bit strictly below position lo
    for (int i = 0; i < a.n; ++i) {
        const int base = 64 * i;
        if (base >= lo) break;
        const uint64_t mask = lo - base >= 64 ? ~0ull : ((1ull << (lo - base)) - 1);
        if (a.w[i] & mask) return true;
    }
    return false;
}

DEC_HD double dec_ldexp_any(uint64_t q, int e2) {  // q * 2^e2, exact (callers keep q within the precision available)
    double r = (double)q;
    while (e2 > 0) { const int s = e2 > 512 ? 512 : e2; r *= dec_ldexp(1, s); e2 -= s; }
    while (e2 < 0) { const int s = e2 < -512 ? 512 : -e2; r *= dec_ldexp(1, -s); e2 += s; }
    return r;
}

// the exact value is (Q + theta) * 2^E, 0 <= theta < 1, sticky = theta > 0, Q has 64 or 65 bits -> nearest-even double
DEC_HD double dec_round_qe(dec_u128 Q, bool sticky, int E) {
    const int nb = dec_bits128(Q);
    const int top = nb - 1 + E;  // unbiased exponent of the value
    if (top > 1023) {
        const uint64_t inf = 0x7ff0000000000000ull;
        return *reinterpret_cast<const double*>(&inf);
    }
    int p = 53;
    if (top < -1022) p = 53 - (-1022 - top);  // subnormal: fewer mantissa bits
    if (p < 0) return 0.0;
    const int shift = nb - p;  // >= 11
    uint64_t q = shift >= 128 ? 0 : (uint64_t)(Q >> shift);
    const dec_u128 rem = shift >= 128 ? Q : (Q & ((((dec_u128)1) << shift) - 1));
    const dec_u128 half = ((dec_u128)1) << (shift - 1);
    if (rem > half || (rem == half && (sticky || (q & 1)))) ++q;
    if (q == 0) return 0.0;
    // q <= 2^53; q * 2^(E + shift) is representable (for subnormals by construction of p), except the carry to inf
    if (E + shift > 971 && q >= (1ull << 53)) {
        const uint64_t inf = 0x7ff0000000000000ull;
        return *reinterpret_cast<const double*>(&inf);
    }
    return dec_ldexp_any(q, E + shift);
}

// M * 10^e10 (M != 0, a multi-word integer of at most ~200 bits) correctly rounded, any e10
DEC_HD double dec_scale_exact_big(DecBig& M, long long e10) {
    const int bm = big_bits(M);
    if (e10 > 310) {
        const uint64_t inf = 0x7ff0000000000000ull;
        return *reinterpret_cast<const double*>(&inf);
    }
    if (e10 < -345 - 61) return 0.0;               // below 10^-345 even with 61 more digits in M
    if (e10 >= 0) {
        big_pow5(M, (int)e10);                     // value = M * 2^e10
        const int nb = big_bits(M);
        const int sh = nb > 64 ? nb - 64 : 0;
        const uint64_t Q = big_extract64(M, sh);
        return dec_round_qe((dec_u128)Q, sh > 0 && big_any_below(M, sh), sh + (int)e10);
    }
    const int k = (int)-e10;
    if (k > 345 + 61) return 0.0;
    DecBig D;
    big_set(D, 1);
    big_pow5(D, k);                                // value = M / D * 2^-k   (5^406 = 943 bits: fits)
    const int bd = big_bits(D);
    // N = M << a with a = bd - bm + 64 (> 0 or < 0): the quotient has 64 or 65 bits.  Shift-subtract division:
    // start the remainder with the top bd - 1 bits of N, then bring down the remaining 66 bits one at a time.
    const int a = bd - bm + 64;
    const int nbits = bm + a;                      // = bd + 64
    DecBig R;
    big_set(R, 0);
    bool dropped = false;                          // a < 0: the low -a bits of M never enter the division
    if (a < 0) dropped = big_any_below(M, -a);
    int i = nbits - 1;
    for (; i > nbits - bd; --i) {                  // bd - 1 bits: no quotient bit can be set yet
        const int mb = i - a;                      // bit of M (negative: the zero padding)
        big_shl1_or(R, mb >= 0 && big_bit(M, mb));
    }
    dec_u128 Q = 0;
    for (; i >= 0; --i) {
        const int mb = i - a;
        big_shl1_or(R, mb >= 0 && big_bit(M, mb));
        Q <<= 1;
        if (big_cmp(R, D) >= 0) { big_sub(R, D); Q |= 1; }
    }
    return dec_round_qe(Q, R.n != 0 || dropped, -a - k);
}
DEC_HD double dec_scale_exact(uint64_t m, long long e10) {
    DecBig M;
    big_set(M, m);
    return dec_scale_exact_big(M, e10);
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
            // only the bare word is the "missing" marker load_csv tests for (v.lower() == "nan",
            // train_tf_ps.py:129); "+nan" / "-nan" go through float() and stay NaN values
            if (i == 0) *flags |= DEC_FLAG_NAN;
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
    const int digits_at = i;  // a numeral with 20..60 significant digits is re-read into a multi-word mantissa
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
    const int digits_end = i;
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
        // 20..60 significant digits: exact through a multi-word mantissa; longer ones are refused
        DecBig M;
        big_set(M, 0);
        int sig2 = 0, frac2 = 0, extra_int = 0;
        bool seen_dot = false, tail2 = false;
        for (int j = digits_at; j < digits_end; ++j) {
            if (p[j] == '.') { seen_dot = true; continue; }
            const unsigned d = p[j] - '0';
            if (sig2 == 0 && d == 0) { frac2 += seen_dot ? 1 : 0; continue; }
            if (sig2 < 60) {
                big_mul_small(M, 10);
                uint64_t carry = d;  // M += d
                for (int w = 0; w < DEC_WORDS && carry; ++w) {
                    const uint64_t t = M.w[w] + carry;
                    carry = t < M.w[w] ? 1 : 0;
                    M.w[w] = t;
                    if (w >= M.n && t) M.n = w + 1;
                }
                ++sig2;
                frac2 += seen_dot ? 1 : 0;
            } else {
                extra_int += seen_dot ? 0 : 1;
                tail2 |= d != 0;
            }
        }
        if (tail2) return DEC_INEXACT;
        r = dec_scale_exact_big(M, e10 - ((long long)dropped - frac_taken) + ((long long)extra_int - frac2));
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
        r = dec_scale_exact(m, e10);
    }
    *out = neg ? -r : r;
    return DEC_OK;
}

}  // namespace etl
