/* cli_writer_check.c - test harness (CPU, no device): the host driver's OWN write_final_state() and
 * write_av_vels() (src/d2q9-bgk.c, compiled in by #include with main renamed) on a synthetic ragged lattice,
 * compared byte for byte with the reference's fprintf formats (d2q9-bgk.c:739, 747).  Run in a scratch
 * directory: it writes final_state.dat / av_vels.dat there.  Exit status 0 = identical. */
#define main lbm_cli_main
#include "../src/d2q9-bgk.c"
#undef main

static unsigned long long rs = 0x9E3779B97F4A7C15ULL;
static unsigned rnd(void) { rs ^= rs << 13; rs ^= rs >> 7; rs ^= rs << 17; return (unsigned)(rs >> 20); }

static float pick(void)
{
    const unsigned r = rnd() % 16u;
    unsigned u;
    float f;
    if (r == 0) return 0.f;
    if (r == 1) return -0.f;
    if (r == 2) { u = rnd(); memcpy(&f, &u, 4); return (f != f || f - f != 0.f) ? 1.5f : f; }   /* any finite bit pattern */
    u = (rnd() & 0x807FFFFFu) | ((90u + rnd() % 38u) << 23);                                  /* 1e-11 .. 2 */
    memcpy(&f, &u, 4);
    return f;
}

static int same_file(const char* name, const char* want, size_t len)
{
    FILE* fp = fopen(name, "rb");
    if (!fp) { printf("cannot open %s\n", name); return 0; }
    char* got = (char*)malloc(len + 16);
    const size_t n = fread(got, 1, len + 16, fp);
    fclose(fp);
    const int ok = (n == len) && memcmp(got, want, len) == 0;
    if (!ok) printf("%s differs (%zu bytes, expected %zu)\n", name, n, len);
    free(got);
    return ok;
}

int main(int argc, char** argv)
{
    run_params p;
    memset(&p, 0, sizeof p);
    p.nx = argc > 1 ? atoi(argv[1]) : 260;
    p.ny = argc > 2 ? atoi(argv[2]) : 37;
    p.max_iters = 1234;
    const size_t n = (size_t)p.nx * p.ny;
    float* f = (float*)malloc(sizeof(float) * 4 * n);
    int* mask = (int*)malloc(sizeof(int) * n);
    float* av = (float*)malloc(sizeof(float) * p.max_iters);
    for (size_t i = 0; i < 4 * n; ++i) f[i] = pick();
    for (size_t i = 0; i < n; ++i) mask[i] = (rnd() % 5u) == 0u;
    for (int i = 0; i < p.max_iters; ++i) av[i] = pick();

    const double t0 = wall();
    write_final_state(&p, mask, f, f + n, f + 2 * n, f + 3 * n);
    const double t1 = wall();
    write_av_vels(&p, av);
    printf("write_final_state: %.3f s for %zu cells\n", t1 - t0, n);
    if (argc > 3) return 0;                    /* timing only */

    char* want = (char*)malloc(n * 128 + (size_t)p.max_iters * 48 + 64);
    char* w = want;
    for (int jj = 0; jj < p.ny; jj++)
        for (int ii = 0; ii < p.nx; ii++) {
            const size_t c = (size_t)jj * p.nx + ii;
            w += sprintf(w, "%d %d %.12E %.12E %.12E %.12E %d\n", ii, jj, f[c], f[n + c], f[2 * n + c], f[3 * n + c], mask[c]);
        }
    int ok = same_file(FINALSTATEFILE, want, (size_t)(w - want));
    w = want;
    for (int ii = 0; ii < p.max_iters; ii++) w += sprintf(w, "%d:\t%.12E\n", ii, av[ii]);
    ok = same_file(AVVELSFILE, want, (size_t)(w - want)) && ok;
    printf("%s\n", ok ? "writers identical to the reference formats" : "MISMATCH");
    return ok ? 0 : 1;
}
