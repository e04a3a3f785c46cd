/* fmt_e12.h - exact, fast text formatting for final_state.dat / av_vels.dat.
 *
 * The reference prints every field with "%.12E" (d2q9-bgk.c:739, 747): 13 significant decimal digits of a
 * float promoted to double, correctly rounded (glibc rounds the EXACT binary value half-to-even).  For the
 * 1.4 GB final_state.dat of 16384x1024 sprintf is the cost of the whole output path, so the common case is
 * done here in integer arithmetic: a float is m * 2^e with m < 2^24, hence for e < 0
 *     digits = round_half_even(m * 10^q / 2^-e),  q = 12 - floor(log10 v),
 * and m * 10^q fits an unsigned __int128 for q <= 31, i.e. for 1e-19 <= |v| < 2^24.  Everything else (zero is
 * handled, larger / smaller magnitudes, inf, nan) goes through snprintf.  tests/fmt_check.c compares the two
 * byte for byte (`fmt_check all`: every one of the 2^32 bit patterns, 0 mismatches; the CPU suite runs a sample).
 */
#ifndef FMT_E12_H
#define FMT_E12_H

#include <stdint.h>
#include <stdio.h>
#include <string.h>

typedef unsigned __int128 fmt_u128;

static const char fmt_digit_pairs[201] =
    "00010203040506070809101112131415161718192021222324252627282930313233343536373839"
    "40414243444546474849505152535455565758596061626364656667686970717273747576777879"
    "8081828384858687888990919293949596979899";

/* 10^q for q = 0..31 */
static const fmt_u128 fmt_pow10_table[32] = {
    1ULL, 10ULL, 100ULL, 1000ULL, 10000ULL, 100000ULL, 1000000ULL, 10000000ULL, 100000000ULL, 1000000000ULL, 10000000000ULL, 100000000000ULL, 1000000000000ULL, 10000000000000ULL, 100000000000000ULL, 1000000000000000ULL, 10000000000000000ULL, 100000000000000000ULL, 1000000000000000000ULL, 10000000000000000000ULL,
    (fmt_u128)10000000000000000000ULL * 10ULL, (fmt_u128)10000000000000000000ULL * 100ULL, (fmt_u128)10000000000000000000ULL * 1000ULL, (fmt_u128)10000000000000000000ULL * 10000ULL, (fmt_u128)10000000000000000000ULL * 100000ULL, (fmt_u128)10000000000000000000ULL * 1000000ULL, (fmt_u128)10000000000000000000ULL * 10000000ULL, (fmt_u128)10000000000000000000ULL * 100000000ULL, (fmt_u128)10000000000000000000ULL * 1000000000ULL, (fmt_u128)10000000000000000000ULL * 10000000000ULL, (fmt_u128)10000000000000000000ULL * 100000000000ULL, (fmt_u128)10000000000000000000ULL * 1000000000000ULL};
static inline fmt_u128 fmt_pow10(int q) { return fmt_pow10_table[q]; }

/* Non-negative int as decimal; returns the advanced pointer. */
static inline char* fmt_uint(char* w, unsigned v)
{
    char tmp[12];
    int n = 0;
    do { tmp[n++] = (char)('0' + v % 10u); v /= 10u; } while (v);
    while (n) *w++ = tmp[--n];
    return w;
}

static inline char* fmt_int(char* w, int v)
{
    if (v < 0) { *w++ = '-'; return fmt_uint(w, 0u - (unsigned)v); }
    return fmt_uint(w, (unsigned)v);
}

/* "%.12E" of a float (as printf sees it: promoted to double); returns the advanced pointer. */
static inline char* fmt_e12(char* w, float f)
{
    uint32_t bits;
    memcpy(&bits, &f, 4);
    const uint32_t mant = bits & 0x7FFFFFu, ex = (bits >> 23) & 0xFFu;
    if (ex == 0xFFu) return w + sprintf(w, "%.12E", (double)f);          /* inf / nan: printf's spelling */
    if (bits & 0x80000000u) *w++ = '-';
    if (ex == 0 && mant == 0) { memcpy(w, "0.000000000000E+00", 18); return w + 18; }
    /* v = m * 2^e2 */
    const uint32_t m = ex ? (mant | 0x800000u) : mant;
    const int e2 = (ex ? (int)ex : 1) - 150;
    const int s = -e2;                                   /* v = m / 2^s */
    int fast = s >= 1 && s <= 100;
    uint64_t D = 0;
    int k = 0;
    if (fast) {
        const int b = (31 - __builtin_clz(m)) - s;       /* floor(log2 v) */
        k = (b * 1233) >> 12;                            /* ~floor(b * log10(2)), off by at most one either way */
        fmt_u128 N = 0, D0 = 0;
        for (int tries = 0; tries < 4; ++tries) {
            const int q = 12 - k;
            if (q < 0 || q > 31) { fast = 0; break; }
            N = (fmt_u128)m * fmt_pow10(q);
            D0 = N >> s;
            if (D0 >= (fmt_u128)10000000000000ULL) { ++k; continue; }
            if (D0 < (fmt_u128)1000000000000ULL) { --k; continue; }
            break;
        }
        if (fast && (D0 >= (fmt_u128)10000000000000ULL || D0 < (fmt_u128)1000000000000ULL)) fast = 0;
        if (fast) {
            const fmt_u128 rem = N & (((fmt_u128)1 << s) - 1), half = (fmt_u128)1 << (s - 1);
            D = (uint64_t)D0;
            if (rem > half || (rem == half && (D & 1u))) ++D;
            if (D == 10000000000000ULL) { D = 1000000000000ULL; ++k; }      /* 9.999...5 rounds up to 1.000...E+(k+1) */
        }
    }
    if (!fast) {
        /* the sign (if any) is already written; print the magnitude */
        union { float f; uint32_t u; } a;
        a.u = bits & 0x7FFFFFFFu;
        return w + sprintf(w, "%.12E", (double)a.f);
    }
    /* D = d.dddddddddddd (13 digits) */
    char dig[13];
    uint64_t r = D;
    for (int i = 11; i >= 1; i -= 2) {
        const unsigned two = (unsigned)(r % 100u);
        r /= 100u;
        dig[i] = fmt_digit_pairs[2 * two];
        dig[i + 1] = fmt_digit_pairs[2 * two + 1];
    }
    dig[0] = (char)('0' + (unsigned)r);
    w[0] = dig[0];
    w[1] = '.';
    memcpy(w + 2, dig + 1, 12);
    w += 14;
    *w++ = 'E';
    unsigned ak;
    if (k < 0) { *w++ = '-'; ak = (unsigned)(-k); } else { *w++ = '+'; ak = (unsigned)k; }
    *w++ = fmt_digit_pairs[2 * ak];                      /* |k| <= 19 on the fast path: always two digits */
    *w++ = fmt_digit_pairs[2 * ak + 1];
    return w;
}

#endif
