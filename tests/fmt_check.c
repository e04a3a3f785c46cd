/* fmt_check.c - test harness (CPU): src/fmt_e12.h against snprintf("%.12E") byte for byte.
 * usage: fmt_check <millions of random bit patterns> ; exit status 0 = all identical.  Prints the first mismatches. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../src/fmt_e12.h"

static unsigned long long bad = 0, total = 0, fast = 0;

static void check(float f)
{
    char a[64], b[64];
    char* e = fmt_e12(a, f);
    *e = 0;
    snprintf(b, sizeof b, "%.12E", (double)f);
    ++total;
    if (strcmp(a, b) != 0) {
        if (bad < 10) { unsigned u; memcpy(&u, &f, 4); printf("MISMATCH bits=%08x  got %s  want %s\n", u, a, b); }
        ++bad;
    }
}

static unsigned long long rng = 88172645463325252ULL;
static unsigned next32(void) { rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17; return (unsigned)(rng >> 16); }

/* fmt_check all : EVERY one of the 2^32 bit patterns (OpenMP; ~6 min on 8 cores). */
static int exhaustive(void)
{
    unsigned long long nbad = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : nbad)
    for (int hi = 0; hi < 65536; ++hi) {
        char a[64], b[64];
        for (unsigned lo = 0; lo < 65536u; ++lo) {
            const unsigned u = ((unsigned)hi << 16) | lo;
            float f;
            memcpy(&f, &u, 4);
            char* e = fmt_e12(a, f);
            *e = 0;
            snprintf(b, sizeof b, "%.12E", (double)f);
            if (strcmp(a, b) != 0) {
                if (nbad < 5) printf("MISMATCH bits=%08x  got %s  want %s\n", u, a, b);
                ++nbad;
            }
        }
    }
    printf("all 4294967296 bit patterns checked, %llu mismatches\n", nbad);
    return nbad ? 1 : 0;
}

int main(int argc, char** argv)
{
    if (argc > 1 && strcmp(argv[1], "all") == 0) return exhaustive();
    const unsigned long long n = (argc > 1 ? strtoull(argv[1], 0, 10) : 20ULL) * 1000000ULL;
    /* 1. special values and every power of two / its neighbours */
    const unsigned specials[] = {0x00000000u, 0x80000000u, 0x00000001u, 0x007FFFFFu, 0x00800000u, 0x7F7FFFFFu, 0x7F800000u,
                                 0xFF800000u, 0x7FC00000u, 0x3F800000u, 0xBF800000u, 0x3DCCCCCDu /* 0.1f */, 0x3D088889u};
    for (unsigned i = 0; i < sizeof specials / sizeof specials[0]; ++i) { float f; memcpy(&f, &specials[i], 4); check(f); }
    for (unsigned ex = 0; ex < 255; ++ex)
        for (int d = -3; d <= 3; ++d) {
            unsigned u = (ex << 23) + (unsigned)d;
            float f; memcpy(&f, &u, 4); check(f);
            u |= 0x80000000u; memcpy(&f, &u, 4); check(f);
        }
    /* 2. values that round up across a power of ten (9.9999999...e-k) and decimal-looking values */
    for (int k = -25; k <= 12; ++k)
        for (int j = -40; j <= 40; ++j) {
            double v = 1.0; for (int i = 0; i < (k < 0 ? -k : k); ++i) v = k < 0 ? v / 10.0 : v * 10.0;
            float f = (float)v; unsigned u; memcpy(&u, &f, 4); u += (unsigned)j; memcpy(&f, &u, 4); check(f);
            f = (float)(v * 0.1234567890123); check(f); check(-f);
        }
    /* 3. random bit patterns: all exponents, then LBM-like magnitudes (1e-9 .. 1) */
    for (unsigned long long i = 0; i < n; ++i) { unsigned u = next32(); float f; memcpy(&f, &u, 4); check(f); }
    for (unsigned long long i = 0; i < n; ++i) {
        unsigned u = (next32() & 0x807FFFFFu) | ((97u + next32() % 31u) << 23);
        float f; memcpy(&f, &u, 4); check(f);
    }
    /* 4. integers */
    const int ints[] = {0, 1, 9, 10, 99, 100, 16383, 16384, 1023, 65535, 2147483647, -1, -2147483647 - 1, 1000000};
    for (unsigned i = 0; i < sizeof ints / sizeof ints[0]; ++i) {
        char a[32], b[32];
        *fmt_int(a, ints[i]) = 0;
        snprintf(b, sizeof b, "%d", ints[i]);
        ++total;
        if (strcmp(a, b)) { printf("MISMATCH int %s vs %s\n", a, b); ++bad; }
    }
    (void)fast;
    printf("%llu values checked, %llu mismatches\n", total, bad);
    return bad ? 1 : 0;
}
