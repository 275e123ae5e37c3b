// dnp_textio.cu -- point-cloud text IO and the server wire format on the device.
//
//   xyz2tensor   util.py:53-69   one point per line, tokens split on single spaces, 3 or 6 columns, lines
//                                containing "nan" skipped; every token goes through Python's float() (correctly
//                                rounded decimal -> double) and torch.tensor (double -> float32, round to nearest).
//   export_pc    util.py:46-51   str(value.item()) per value = Python's repr of the float32 widened to double:
//                                the SHORTEST decimal that reads back as that double, joined by ' ' and '\n'.
//   wire format  socket_server_para.py:145-166,191: request payload = N x 3 float64, reply = N x 6 float64.
//
// Both directions are per-line / per-value independent byte work: thread per line (parse) or per value (format),
// HBM bound.  Exactness comes from small fixed-size big integers (16 x 32-bit limbs, in registers / local memory):
//   * parse:  value = w * 10^q with w the first 38 significant digits.  q >= 0: the product, q < 0: floor division
//             of w << s by 10^-q with a sticky bit (nested floor divisions by 10^9 compose exactly), then one
//             round-half-even to 53 bits.  More than 38 digits: accepted when w and w + 1 round to the same double.
//   * format: Steele-White / Burger-Dybvig free-format digit generation with IEEE boundary rules (boundaries
//             included for an even mantissa, lower gap halved at a power of two), ties in the last digit to even
//             as David Gay's dtoa (CPython's repr) does, then CPython's float_repr_style layout ('1e-05', '0.001',
//             '1.5e+16', '12345.0', 'inf', 'nan', '-0.0').
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>

#include "dnp_common.cuh"
#include "dnp_pairs.h"
#include "dnp_scores.h"
#include "dipole.h"

namespace dnp {
namespace {

// ------------------------------------------------------------------ fixed-size big integers
constexpr int kLimbs = 16;                 // 512 bits
struct Big {
    uint32_t d[kLimbs];
};

__device__ __forceinline__ void big_zero(Big& a) {
#pragma unroll
    for (int i = 0; i < kLimbs; ++i) a.d[i] = 0;
}
__device__ __forceinline__ void big_set64(Big& a, unsigned long long v) {
    big_zero(a);
    a.d[0] = (uint32_t)v;
    a.d[1] = (uint32_t)(v >> 32);
}
__device__ __forceinline__ void big_mul_small(Big& a, uint32_t m) {
    unsigned long long carry = 0;
#pragma unroll
    for (int i = 0; i < kLimbs; ++i) {
        const unsigned long long t = (unsigned long long)a.d[i] * m + carry;
        a.d[i] = (uint32_t)t;
        carry = t >> 32;
    }
}
__device__ __forceinline__ void big_mul_pow10(Big& a, int k) {
    for (; k >= 9; k -= 9) big_mul_small(a, 1000000000u);
    uint32_t m = 1;
    for (int i = 0; i < k; ++i) m *= 10u;
    if (m > 1) big_mul_small(a, m);
}
__device__ __forceinline__ void big_shl(Big& a, int bits) {       // bits < 32 * kLimbs
    const int limbs = bits >> 5, sh = bits & 31;
    for (int i = kLimbs - 1; i >= 0; --i) {
        const int j = i - limbs;
        uint32_t v = 0;
        if (j >= 0) {
            v = a.d[j] << sh;
            if (sh && j >= 1) v |= a.d[j - 1] >> (32 - sh);
        }
        a.d[i] = v;
    }
}
__device__ __forceinline__ int big_cmp(const Big& a, const Big& b) {
    for (int i = kLimbs - 1; i >= 0; --i) {
        if (a.d[i] != b.d[i]) return a.d[i] > b.d[i] ? 1 : -1;
    }
    return 0;
}
__device__ __forceinline__ void big_add(Big& a, const Big& b) {
    unsigned long long carry = 0;
#pragma unroll
    for (int i = 0; i < kLimbs; ++i) {
        const unsigned long long t = (unsigned long long)a.d[i] + b.d[i] + carry;
        a.d[i] = (uint32_t)t;
        carry = t >> 32;
    }
}
// The same operations on the low `nl` limbs only (the callers know how wide their numbers are): the digit loops
// run on 3-6 limbs for ordinary values instead of 16.
__device__ __forceinline__ void big_mul_small_n(Big& a, uint32_t m, int nl) {
    unsigned long long carry = 0;
    for (int i = 0; i < nl; ++i) {
        const unsigned long long t = (unsigned long long)a.d[i] * m + carry;
        a.d[i] = (uint32_t)t;
        carry = t >> 32;
    }
}
__device__ __forceinline__ bool big_div_1e9_n(Big& a, int nl) {    // a = floor(a / 10^9); true if a remainder is left
    unsigned long long rem = 0;
    for (int i = nl - 1; i >= 0; --i) {
        const unsigned long long t = (rem << 32) | a.d[i];
        const unsigned long long q = t / 1000000000ull;           // constant divisor: multiply-high, no division loop
        a.d[i] = (uint32_t)q;
        rem = t - q * 1000000000ull;
    }
    return rem != 0;
}
__device__ __forceinline__ int big_cmp_n(const Big& a, const Big& b, int nl) {
    for (int i = nl - 1; i >= 0; --i) {
        if (a.d[i] != b.d[i]) return a.d[i] > b.d[i] ? 1 : -1;
    }
    return 0;
}
__device__ __forceinline__ void big_add_n(Big& a, const Big& b, int nl) {
    unsigned long long carry = 0;
    for (int i = 0; i < nl; ++i) {
        const unsigned long long t = (unsigned long long)a.d[i] + b.d[i] + carry;
        a.d[i] = (uint32_t)t;
        carry = t >> 32;
    }
}
__device__ __forceinline__ void big_sub_n(Big& a, const Big& b, int nl) {   // a >= b
    unsigned int borrow = 0;
    for (int i = 0; i < nl; ++i) {
        const unsigned long long t = (unsigned long long)a.d[i] - b.d[i] - borrow;
        a.d[i] = (uint32_t)t;
        borrow = (unsigned int)(t >> 63);
    }
}
__device__ __forceinline__ int big_top_bit(const Big& a) {        // index of the highest set bit, -1 for zero
    for (int i = kLimbs - 1; i >= 0; --i)
        if (a.d[i]) return i * 32 + 31 - __clz(a.d[i]);
    return -1;
}
__device__ __forceinline__ int big_bit(const Big& a, int pos) { return (a.d[pos >> 5] >> (pos & 31)) & 1u; }
__device__ __forceinline__ bool big_any_below(const Big& a, int pos) {   // any set bit at index < pos
    const int limb = pos >> 5;
    for (int i = 0; i < limb; ++i)
        if (a.d[i]) return true;
    const int sh = pos & 31;
    return sh ? (a.d[limb] & ((1u << sh) - 1u)) != 0 : false;
}

// ------------------------------------------------------------------ decimal -> double, correctly rounded
// top 53 bits of `a` (value a * 2^exp2 plus a sticky tail) rounded half-to-even
__device__ double big_round_to_double(const Big& a, int exp2, bool sticky) {
    const int h = big_top_bit(a);
    if (h < 0) return 0.0;
    if (h <= 52) {
        const unsigned long long m = ((unsigned long long)a.d[1] << 32) | a.d[0];
        return ldexp((double)m, exp2);                         // exact (callers guarantee no sticky tail here)
    }
    const int lo = h - 52;
    const int idx = lo >> 5, sh = lo & 31;
    unsigned long long v = a.d[idx];
    if (idx + 1 < kLimbs) v |= (unsigned long long)a.d[idx + 1] << 32;
    unsigned long long m = v >> sh;
    if (sh && idx + 2 < kLimbs) m |= (unsigned long long)a.d[idx + 2] << (64 - sh);
    m &= (1ull << 53) - 1ull;
    const int guard = big_bit(a, lo - 1);
    const bool rest = sticky || big_any_below(a, lo - 1);
    int e = lo + exp2;
    if (guard && (rest || (m & 1ull))) {
        ++m;
        if (m == (1ull << 53)) { m >>= 1; ++e; }
    }
    return ldexp((double)m, e);
}

constexpr int kMaxPow10 = 60;      // w * 10^q, q in [-kMinPow10, kMaxPow10], w < 10^38 < 2^127: fits 512 bits
constexpr int kMinPow10 = 76;

__constant__ double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                  1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

__device__ double dec_to_double(Big a, int q) {
    if (q >= 0) {
        big_mul_pow10(a, q);
        return big_round_to_double(a, 0, false);
    }
    // w / 10^n = (w * 10^pad) / 10^(n + pad): every division is by the constant 10^9
    const int n = -q;
    const int pad = (9 - n % 9) % 9;
    big_mul_pow10(a, pad);
    const int s = 32 * ((57 + (n * 3402) / 1024 + 1 + 31) / 32);   // quotient keeps >= 56 significant bits
    big_shl(a, s);
    const int nl = min(kLimbs, big_top_bit(a) / 32 + 1);
    bool sticky = false;
    for (int k = n + pad; k > 0; k -= 9) sticky |= big_div_1e9_n(a, nl);
    return big_round_to_double(a, -s, sticky);
}

// w < 2^53 and |q| <= 22: both operands are exact doubles, one IEEE multiply / divide is correctly rounded
__device__ __forceinline__ bool dec_fast_path(unsigned long long w, int q, double& out) {
    if (w >> 53) return false;
    if (q >= 0 && q <= 22) { out = __dmul_rn((double)w, kPow10[q]); return true; }
    if (q < 0 && q >= -22) { out = __ddiv_rn((double)w, kPow10[-q]); return true; }
    return false;
}

__device__ __forceinline__ bool is_py_space(unsigned char c) {     // str.strip() / float() whitespace, ASCII part
    return c == ' ' || (c >= 9 && c <= 13) || (c >= 0x1c && c <= 0x1f);
}
__device__ __forceinline__ unsigned char lower(unsigned char c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; }

__device__ bool match_word(const char* s, int len, const char* word, int wlen) {
    if (len != wlen) return false;
    for (int i = 0; i < len; ++i)
        if (lower((unsigned char)s[i]) != (unsigned char)word[i]) return false;
    return true;
}

// Python float(token) for the literals a point-cloud file holds.  false = not a literal this parser resolves
// (malformed, underscores, or a decimal exponent outside [-76, 60] after normalisation).
__device__ bool parse_float_token(const char* s, int len, double& out) {
    while (len > 0 && is_py_space((unsigned char)s[0])) { ++s; --len; }
    while (len > 0 && is_py_space((unsigned char)s[len - 1])) --len;
    if (len == 0) return false;
    bool neg = false;
    if (s[0] == '+' || s[0] == '-') { neg = s[0] == '-'; ++s; --len; }
    if (len == 0) return false;
    if (match_word(s, len, "inf", 3) || match_word(s, len, "infinity", 8)) {
        out = neg ? -INFINITY : INFINITY;
        return true;
    }
    if (match_word(s, len, "nan", 3)) {
        out = neg ? -NAN : NAN;
        return true;
    }
    unsigned long long w = 0, w2 = 0;                          // digits 1..19 and 20..38
    int nd = 0, frac_shift = 0, int_extra = 0, i = 0;
    bool any_digit = false, dp_seen = false, trunc = false;
    for (; i < len; ++i) {
        const unsigned char c = (unsigned char)s[i];
        if (c >= '0' && c <= '9') {
            any_digit = true;
            const int dgt = c - '0';
            if (nd == 0 && dgt == 0) {                         // leading zero: only moves the point
                if (dp_seen) ++frac_shift;
                continue;
            }
            if (nd < 19) {
                w = w * 10ull + (unsigned long long)dgt;
                ++nd;
                if (dp_seen) ++frac_shift;
            } else if (nd < 38) {
                w2 = w2 * 10ull + (unsigned long long)dgt;
                ++nd;
                if (dp_seen) ++frac_shift;
            } else {
                if (dgt) trunc = true;
                if (!dp_seen) ++int_extra;
            }
        } else if (c == '.') {
            if (dp_seen) return false;
            dp_seen = true;
        } else if (c == 'e' || c == 'E') {
            break;
        } else {
            return false;
        }
    }
    if (!any_digit) return false;
    int ex = 0;
    if (i < len) {                                             // exponent part
        ++i;
        bool eneg = false;
        if (i < len && (s[i] == '+' || s[i] == '-')) { eneg = s[i] == '-'; ++i; }
        if (i >= len) return false;
        for (; i < len; ++i) {
            const unsigned char c = (unsigned char)s[i];
            if (c < '0' || c > '9') return false;
            if (ex < 100000) ex = ex * 10 + (c - '0');
        }
        if (eneg) ex = -ex;
    }
    if (w == 0) {
        out = neg ? -0.0 : 0.0;
        return true;
    }
    const int q = ex + int_extra - frac_shift;
    if (q > kMaxPow10 || q < -kMinPow10) return false;
    double fast;
    if (nd <= 19 && !trunc && dec_fast_path(w, q, fast)) {
        out = neg ? -fast : fast;
        return true;
    }
    Big a;
    big_set64(a, w);
    if (nd > 19) {
        big_mul_pow10(a, nd - 19);
        Big lo;
        big_set64(lo, w2);
        big_add(a, lo);
    }
    double v = dec_to_double(a, q);
    if (trunc) {                                               // true value lies in (w, w + 1) * 10^q
        Big one;
        big_set64(one, 1ull);
        big_add(a, one);
        if (dec_to_double(a, q) != v) return false;
    }
    out = neg ? -v : v;
    return true;
}

// ------------------------------------------------------------------ parse kernels
struct IsNewline {
    const char* txt;
    __device__ bool operator()(int64_t i) const { return txt[i] == '\n'; }
};

__device__ __forceinline__ void line_span(const int64_t* nl_pos, int64_t n_nl, int64_t n_bytes, int64_t k, int64_t& a,
                                          int64_t& b) {
    a = (k == 0) ? 0 : nl_pos[k - 1] + 1;
    b = (k < n_nl) ? nl_pos[k] : n_bytes;
}

// per line: strip, skip "nan" lines, count the single-space separated tokens -> 0 / 3 / 6 columns
__global__ void classify_lines_kernel(const char* __restrict__ txt, int64_t n_bytes, const int64_t* __restrict__ nl_pos,
                                      const int64_t* __restrict__ n_nl_ptr, int64_t max_lines,
                                      int32_t* __restrict__ valid, uint8_t* __restrict__ cols) {
    const int64_t n_nl = *n_nl_ptr;
    const int64_t n_lines = min(n_nl + 1, max_lines);
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < max_lines; k += (int64_t)gridDim.x * blockDim.x) {
        int c = 0;
        if (k < n_lines) {
            int64_t a, b;
            line_span(nl_pos, n_nl, n_bytes, k, a, b);
            while (a < b && is_py_space((unsigned char)txt[a])) ++a;
            while (b > a && is_py_space((unsigned char)txt[b - 1])) --b;
            int tokens = 1;
            bool has_nan = false;
            for (int64_t i = a; i < b; ++i) {
                const char ch = txt[i];
                tokens += ch == ' ';
                if (ch == 'n' && i >= a + 2 && txt[i - 1] == 'a' && txt[i - 2] == 'n') has_nan = true;
            }
            if (!has_nan && (tokens == 3 || tokens == 6)) c = tokens;
        }
        cols[k] = (uint8_t)c;
        valid[k] = c != 0;
    }
}

__global__ void parse_lines_kernel(const char* __restrict__ txt, int64_t n_bytes, const int64_t* __restrict__ nl_pos,
                                   const int64_t* __restrict__ n_nl_ptr, int64_t max_lines,
                                   const int32_t* __restrict__ row_of, const uint8_t* __restrict__ cols,
                                   float* __restrict__ out, int64_t max_rows, unsigned long long* __restrict__ counts) {
    const int64_t n_nl = *n_nl_ptr;
    const int64_t n_lines = min(n_nl + 1, max_lines);
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n_lines; k += (int64_t)gridDim.x * blockDim.x) {
        const int c = cols[k];
        if (!c) continue;
        const int64_t row = row_of[k];
        if (row >= max_rows) continue;
        int64_t a, b;
        line_span(nl_pos, n_nl, n_bytes, k, a, b);
        while (a < b && is_py_space((unsigned char)txt[a])) ++a;
        while (b > a && is_py_space((unsigned char)txt[b - 1])) --b;
        float* o = out + row * 6;
        int col = 0;
        bool ok = true;
        int64_t t0 = a;
        for (int64_t i = a; i <= b; ++i) {
            if (i == b || txt[i] == ' ') {
                double v = 0.0;
                ok &= parse_float_token(txt + t0, (int)(i - t0), v);
                if (col < 6) o[col] = __double2float_rn(v);
                ++col;
                t0 = i + 1;
            }
        }
        for (; col < 6; ++col) o[col] = 0.f;
        atomicOr(counts + 1, c == 3 ? 1ull : 2ull);
        if (!ok) atomicMin(counts + 2, (unsigned long long)k);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) counts[3] = (unsigned long long)(n_nl + 1);
}

__global__ void finish_parse_kernel(const int32_t* __restrict__ valid, const int32_t* __restrict__ row_of,
                                    int64_t max_lines, unsigned long long* __restrict__ counts) {
    counts[0] = (unsigned long long)(row_of[max_lines - 1] + valid[max_lines - 1]);
}

// ------------------------------------------------------------------ double -> shortest decimal (CPython repr)
// Free-format digit generation (Burger & Dybvig) on 384-bit integers.  digits[0..nd) and decpt with
// v = 0.d1 d2 ... x 10^decpt.  v finite, > 0, of float32 range (|exponent| small enough for 384 bits).
__device__ int shortest_digits(double v, unsigned char* digits, int& decpt) {
    const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
    const int be = (int)((bits >> 52) & 0x7ff);
    const unsigned long long frac = bits & ((1ull << 52) - 1ull);
    const unsigned long long f = be ? (frac | (1ull << 52)) : frac;
    const int e = (be ? be : 1) - 1075;
    const bool even = !(f & 1ull);
    const bool pow2 = (frac == 0) && be > 1;                    // lower neighbour is half as far away
    Big r, s, mp, mm;
    big_set64(r, f);
    big_set64(s, 1);
    big_set64(mp, 1);
    big_set64(mm, 1);
    if (e >= 0) {
        big_shl(r, e + (pow2 ? 2 : 1));
        big_shl(s, pow2 ? 2 : 1);
        big_shl(mp, e + (pow2 ? 1 : 0));
        big_shl(mm, e);
    } else {
        big_shl(r, pow2 ? 2 : 1);
        big_shl(s, -e + (pow2 ? 2 : 1));
        if (pow2) big_shl(mp, 1);
    }
    int k = (int)ceil(log10(v) - 1e-10);
    if (k >= 0) {
        big_mul_pow10(s, k);
    } else {
        big_mul_pow10(r, -k);
        big_mul_pow10(mp, -k);
        big_mul_pow10(mm, -k);
    }
    {
        Big high = r;
        big_add(high, mp);
        const int c = big_cmp(high, s);
        if (even ? c >= 0 : c > 0) {
            ++k;
        } else {
            big_mul_small(r, 10u);
            big_mul_small(mp, 10u);
            big_mul_small(mm, 10u);
        }
    }
    decpt = k;
    // from here on r < 10 s and m+, m- <= 10 s: everything fits in the limbs of s plus one
    const int nl = min(kLimbs, big_top_bit(s) / 32 + 2);
    int nd = 0;
    while (true) {
        int d = 0;
        while (big_cmp_n(r, s, nl) >= 0) { big_sub_n(r, s, nl); ++d; }      // quotient < 10
        const int c1 = big_cmp_n(r, mm, nl);
        const bool tc1 = even ? c1 <= 0 : c1 < 0;
        Big high = r;
        big_add_n(high, mp, nl);
        const int c2 = big_cmp_n(high, s, nl);
        const bool tc2 = even ? c2 >= 0 : c2 > 0;
        if (!tc1 && !tc2 && nd < 24) {
            digits[nd++] = (unsigned char)d;
            big_mul_small_n(r, 10u, nl);
            big_mul_small_n(mp, 10u, nl);
            big_mul_small_n(mm, 10u, nl);
            continue;
        }
        bool up;
        if (tc1 && !tc2) up = false;
        else if (!tc1 && tc2) up = true;
        else {                                                  // both: the digit closer to v, ties to even (dtoa)
            Big twice = r;
            big_add_n(twice, r, nl);
            const int c = big_cmp_n(twice, s, nl);
            up = c > 0 || (c == 0 && (d & 1));
        }
        digits[nd++] = (unsigned char)(d + (up ? 1 : 0));
        break;
    }
    // a last digit of 10 carries
    for (int i = nd - 1; i >= 0 && digits[i] == 10; --i) {
        digits[i] = 0;
        if (i == 0) {
            for (int j = nd - 1; j > 0; --j) digits[j] = digits[j - 1];
            digits[0] = 1;
            ++decpt;
        } else {
            ++digits[i - 1];
        }
    }
    while (nd > 1 && digits[nd - 1] == 0) --nd;                 // trailing zeros carry no information
    return nd;
}

// str(float) of CPython (float_repr_style 'short', format code 'r'): returns the length written to buf (<= 25)
__device__ int format_repr(float x, char* buf) {
    int n = 0;
    const double v = (double)x;
    if (isnan(v)) { buf[0] = 'n'; buf[1] = 'a'; buf[2] = 'n'; return 3; }
    if (signbit(v)) buf[n++] = '-';
    if (isinf(v)) { buf[n++] = 'i'; buf[n++] = 'n'; buf[n++] = 'f'; return n; }
    if (v == 0.0) { buf[n++] = '0'; buf[n++] = '.'; buf[n++] = '0'; return n; }
    unsigned char dg[28];
    int decpt;
    const int nd = shortest_digits(fabs(v), dg, decpt);
    if (decpt <= -4 || decpt > 16) {                            // exponent notation
        buf[n++] = (char)('0' + dg[0]);
        if (nd > 1) {
            buf[n++] = '.';
            for (int i = 1; i < nd; ++i) buf[n++] = (char)('0' + dg[i]);
        }
        buf[n++] = 'e';
        int ex = decpt - 1;
        buf[n++] = ex < 0 ? '-' : '+';
        if (ex < 0) ex = -ex;
        if (ex >= 100) buf[n++] = (char)('0' + ex / 100);
        buf[n++] = (char)('0' + (ex / 10) % 10);
        buf[n++] = (char)('0' + ex % 10);
    } else if (decpt <= 0) {
        buf[n++] = '0';
        buf[n++] = '.';
        for (int i = 0; i < -decpt; ++i) buf[n++] = '0';
        for (int i = 0; i < nd; ++i) buf[n++] = (char)('0' + dg[i]);
    } else {
        for (int i = 0; i < decpt; ++i) buf[n++] = i < nd ? (char)('0' + dg[i]) : '0';
        buf[n++] = '.';
        if (nd > decpt) {
            for (int i = decpt; i < nd; ++i) buf[n++] = (char)('0' + dg[i]);
        } else {
            buf[n++] = '0';
        }
    }
    return n;
}

constexpr int kSlot = 32;          // bytes reserved per formatted value (longest repr of a float32: 24)

// pass 1: one value per thread -> its characters in a fixed slot, its length (+ separator) for the scan
__global__ void format_values_kernel(const float* __restrict__ pts, int64_t n_rows, int cols, int64_t ld_row,
                                     int64_t ld_col, char* __restrict__ slots, int64_t* __restrict__ lens) {
    const int64_t total = n_rows * cols;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / cols;
        const int col = (int)(i - row * cols);
        __align__(16) char buf[kSlot];
        const int n = format_repr(pts[row * ld_row + col * ld_col], buf);
        uint4* dst = reinterpret_cast<uint4*>(slots + i * kSlot);
        const uint4* src = reinterpret_cast<const uint4*>(buf);
        dst[0] = src[0];
        dst[1] = src[1];
        lens[i] = n + (i + 1 < total ? 1 : 0);                  // ' ' inside a row, '\n' between rows, nothing at the end
    }
}

// pass 2: characters to their final position (offsets = exclusive scan of lens)
__global__ void format_scatter_kernel(const char* __restrict__ slots, const int64_t* __restrict__ lens,
                                      const int64_t* __restrict__ offsets, int64_t total, int cols,
                                      char* __restrict__ out, int64_t capacity, int64_t* __restrict__ n_bytes_out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t off = offsets[i];
        const bool last = i + 1 == total;
        const int n = (int)lens[i] - (last ? 0 : 1);
        if (off + lens[i] <= capacity) {
            const char* src = slots + i * kSlot;
            for (int k = 0; k < n; ++k) out[off + k] = src[k];
            if (!last) out[off + n] = ((i + 1) % cols == 0) ? '\n' : ' ';
        }
        if (last) *n_bytes_out = off + n;
    }
}

// ------------------------------------------------------------------ wire format (socket_server_para.py:145-166,191)
// request: N x 3 float64 -> [N, 6] float32 rows with zero normals (util.npxyz2tensor + the float32 cast of the kernels)
__global__ void wire_decode_kernel(const double* __restrict__ xyz, int64_t n, float* __restrict__ out, int64_t ld) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float* o = out + i * ld;
        o[0] = __double2float_rn(xyz[3 * i]);
        o[1] = __double2float_rn(xyz[3 * i + 1]);
        o[2] = __double2float_rn(xyz[3 * i + 2]);
        o[3] = 0.f; o[4] = 0.f; o[5] = 0.f;
    }
}
// reply: [N, 6] float32 rows -> N x 6 float64 (result.astype(np.float64).tobytes())
__global__ void wire_encode_kernel(const float* __restrict__ pts, int64_t n, int64_t ld, double* __restrict__ out) {
    const int64_t total = n * 6;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / 6;
        out[i] = (double)pts[row * ld + (i - row * 6)];
    }
}

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

struct ParseLayout {
    size_t nl_pos, n_nl, valid, row_of, cols, cub_temp, cub_bytes, total;
};

ParseLayout parse_layout(int64_t n_bytes, int64_t max_lines) {
    ParseLayout L{};
    const int nb = (int)(n_bytes < 1 ? 1 : n_bytes);
    const int ml = (int)(max_lines < 1 ? 1 : max_lines);
    size_t sel = 0, scan = 0;
    cub::DeviceSelect::If(nullptr, sel, thrust::counting_iterator<int64_t>(0), (int64_t*)nullptr, (int64_t*)nullptr, nb,
                          IsNewline{nullptr});
    cub::DeviceScan::ExclusiveSum(nullptr, scan, (const int32_t*)nullptr, (int32_t*)nullptr, ml);
    L.cub_bytes = sel > scan ? sel : scan;
    size_t off = 0;
    L.nl_pos = off; off += up256((size_t)ml * 8 + 8);
    L.n_nl = off; off += 256;
    L.valid = off; off += up256((size_t)ml * 4);
    L.row_of = off; off += up256((size_t)ml * 4);
    L.cols = off; off += up256((size_t)ml);
    L.cub_temp = off; off += up256(L.cub_bytes);
    L.total = off;
    return L;
}

struct FormatLayout {
    size_t slots, lens, offsets, cub_temp, cub_bytes, total;
};

FormatLayout format_layout(int64_t n_values) {
    FormatLayout L{};
    const int nv = (int)(n_values < 1 ? 1 : n_values);
    size_t scan = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan, (const int64_t*)nullptr, (int64_t*)nullptr, nv);
    L.cub_bytes = scan;
    size_t off = 0;
    L.slots = off; off += up256((size_t)nv * kSlot);
    L.lens = off; off += up256((size_t)nv * 8);
    L.offsets = off; off += up256((size_t)nv * 8);
    L.cub_temp = off; off += up256(scan);
    L.total = off;
    return L;
}

constexpr int64_t kMaxTextBytes = 0x7fffffff;     // one call: < 2 GiB of text, < 2^31 values

}  // namespace

size_t xyz_parse_workspace_bytes(int64_t n_bytes, int64_t max_lines) {
    if (n_bytes < 0 || n_bytes >= kMaxTextBytes || max_lines < 1 || max_lines >= kMaxTextBytes) return 0;
    return parse_layout(n_bytes, max_lines).total;
}

int run_xyz_parse(const char* txt, int64_t n_bytes, int64_t max_lines, float* out, int64_t max_rows, int64_t* counts,
                  void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    if (n_bytes >= kMaxTextBytes || max_lines >= kMaxTextBytes) {
        set_error("xyz parse: at most 2 GiB of text per call");
        return DNP_ERR_INVALID;
    }
    const ParseLayout L = parse_layout(n_bytes, max_lines);
    if (!workspace || workspace_bytes < L.total) {
        set_error("workspace too small: need %zu bytes (dnp_xyz_parse_workspace_bytes), got %zu", L.total, workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
    int64_t* nl_pos = reinterpret_cast<int64_t*>(ws + L.nl_pos);
    int64_t* n_nl = reinterpret_cast<int64_t*>(ws + L.n_nl);
    int32_t* valid = reinterpret_cast<int32_t*>(ws + L.valid);
    int32_t* row_of = reinterpret_cast<int32_t*>(ws + L.row_of);
    uint8_t* cols = ws + L.cols;
    void* cub_temp = ws + L.cub_temp;
    // counts: rows, column mask (1: 3-column lines seen, 2: 6-column lines seen), first unparsable line, lines
    const unsigned long long init[4] = {0ull, 0ull, ~0ull, 0ull};
    DNP_CHECK_CUDA(cudaMemcpyAsync(counts, init, sizeof(init), cudaMemcpyHostToDevice, stream));
    size_t cub_bytes = L.cub_bytes;
    if (n_bytes > 0) {
        DNP_CHECK_CUDA(cub::DeviceSelect::If(cub_temp, cub_bytes, thrust::counting_iterator<int64_t>(0), nl_pos, n_nl,
                                             (int)n_bytes, IsNewline{txt}, stream));
    } else {
        DNP_CHECK_CUDA(cudaMemsetAsync(n_nl, 0, sizeof(int64_t), stream));
    }
    const int blocks = 148 * 8;
    classify_lines_kernel<<<blocks, 256, 0, stream>>>(txt, n_bytes, nl_pos, n_nl, max_lines, valid, cols);
    cub_bytes = L.cub_bytes;
    DNP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(cub_temp, cub_bytes, valid, row_of, (int)max_lines, stream));
    parse_lines_kernel<<<blocks, 256, 0, stream>>>(txt, n_bytes, nl_pos, n_nl, max_lines, row_of, cols, out, max_rows,
                                                  reinterpret_cast<unsigned long long*>(counts));
    finish_parse_kernel<<<1, 1, 0, stream>>>(valid, row_of, max_lines, reinterpret_cast<unsigned long long*>(counts));
    count_launch(3);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

size_t xyz_format_workspace_bytes(int64_t n_values) {
    if (n_values < 0 || n_values >= kMaxTextBytes / kSlot * 16) return 0;
    return format_layout(n_values).total;
}

int run_xyz_format(const float* pts, int64_t n_rows, int cols, int64_t ld_row, int64_t ld_col, char* out,
                   int64_t capacity, int64_t* n_bytes_out, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    const int64_t total = n_rows * cols;
    if (total == 0) {
        DNP_CHECK_CUDA(cudaMemsetAsync(n_bytes_out, 0, sizeof(int64_t), stream));
        return DNP_OK;
    }
    if (total >= kMaxTextBytes / kSlot * 16) {
        set_error("xyz format: too many values for one call");
        return DNP_ERR_INVALID;
    }
    const FormatLayout L = format_layout(total);
    if (!workspace || workspace_bytes < L.total) {
        set_error("workspace too small: need %zu bytes (dnp_xyz_format_workspace_bytes), got %zu", L.total, workspace_bytes);
        return DNP_ERR_WORKSPACE;
    }
    unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
    char* slots = reinterpret_cast<char*>(ws + L.slots);
    int64_t* lens = reinterpret_cast<int64_t*>(ws + L.lens);
    int64_t* offsets = reinterpret_cast<int64_t*>(ws + L.offsets);
    const int blocks = 148 * 8;
    format_values_kernel<<<blocks, 128, 0, stream>>>(pts, n_rows, cols, ld_row, ld_col, slots, lens);
    size_t cub_bytes = L.cub_bytes;
    DNP_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(ws + L.cub_temp, cub_bytes, lens, offsets, (int)total, stream));
    format_scatter_kernel<<<blocks, 256, 0, stream>>>(slots, lens, offsets, total, cols, out, capacity, n_bytes_out);
    count_launch(2);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

int run_wire_decode(const double* xyz, int64_t n, float* out, int64_t ld, cudaStream_t stream) {
    if (n <= 0) return DNP_OK;
    wire_decode_kernel<<<148 * 4, 256, 0, stream>>>(xyz, n, out, ld);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

int run_wire_encode(const float* pts, int64_t n, int64_t ld, double* out, cudaStream_t stream) {
    if (n <= 0) return DNP_OK;
    wire_encode_kernel<<<148 * 4, 256, 0, stream>>>(pts, n, ld, out);
    count_launch(1);
    DNP_CHECK_CUDA(cudaGetLastError());
    return DNP_OK;
}

}  // namespace dnp
