/*
 * load_generator.c -- `emb_host`: C microbenchmark of the lookup path, plain C against the
 * ABI in include/emb_b200.h.
 *
 * Counterpart of upmem/src/load_generator.c:27-137 (synthetic tables, uniform indices,
 * fixed pooling, 100 timed lookup() calls), with the reference's loose ends closed
 * (SURVEY.md Q8): every column is populated, indices stay inside [0, rows), the result
 * buffers match the shape lookup() writes, and the CPU re-computation it carries
 * (validate_result, :40-65, call commented out there) runs -- here as an exact check of the
 * int32 wrap-sums, not a 1e-6 tolerance.
 *
 *   emb_host [--tables T] [--cols D] [--rows R] [--batches B] [--pooling P] [--iters N]
 *            [--seed S] [--no-validate]
 * Defaults: T, D, B, P from NR_TABLES / NR_COLS / MAX_NR_BATCHES / MAX_INDICES_PER_BATCH
 * (the names upmem/run.sh exports), else the reference's `toy` preset 9 x 64, 64 bags x 32.
 * Prints one JSON line.
 */
#include "emb_b200.h"

#include <dlfcn.h>
#include <libgen.h>
#include <limits.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

/* The library is loaded by path like the reference's callers do (CDLL("./emblib.so"),
 * dlrm_dpu_pytorch.py:123-124): next to this binary, or $EMBLIB_SO.  (Its soname is
 * `emb_host`, upmem/Makefile:54 -- the name of this binary -- so a DT_NEEDED link would
 * resolve to the executable itself.) */
static struct dpu_set_t *(*p_populate_mram)(uint32_t, uint64_t, uint32_t, int32_t *, dpu_runtime_totals *);
static int32_t *(*p_lookup)(uint32_t **, uint32_t **, float **, void *, int64_t);
static int (*p_emb_configure)(uint32_t, uint32_t, uint32_t, uint32_t, int);
static const char *(*p_emb_last_error)(void);
static int (*p_emb_last_timings)(emb_timings *);
static void (*p_emb_shutdown)(void);
#define populate_mram p_populate_mram
#define lookup p_lookup
#define emb_configure p_emb_configure
#define emb_last_error p_emb_last_error
#define emb_last_timings p_emb_last_timings
#define emb_shutdown p_emb_shutdown

static void load_library(const char *argv0)
{
    char path[PATH_MAX], self[PATH_MAX];
    const char *env = getenv("EMBLIB_SO");
    if (env && *env) {
        snprintf(path, sizeof path, "%s", env);
    } else {
        ssize_t n = readlink("/proc/self/exe", self, sizeof self - 1);
        if (n <= 0)
            snprintf(self, sizeof self, "%s", argv0);
        else
            self[n] = 0;
        snprintf(path, sizeof path, "%s/emblib.so", dirname(self));
    }
    void *h = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        fprintf(stderr, "emb_host: %s\n", dlerror());
        exit(1);
    }
#define SYM(name)                                                      \
    do {                                                               \
        *(void **)(&p_##name) = dlsym(h, #name);                       \
        if (!p_##name) {                                               \
            fprintf(stderr, "emb_host: %s lacks %s\n", path, #name);   \
            exit(1);                                                   \
        }                                                              \
    } while (0)
    SYM(populate_mram);
    SYM(lookup);
    SYM(emb_configure);
    SYM(emb_last_error);
    SYM(emb_last_timings);
    SYM(emb_shutdown);
#undef SYM
}

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;

static uint64_t rng_next(void) /* splitmix64 */
{
    uint64_t z = (rng_state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static double now_ms(void)
{
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return t.tv_sec * 1e3 + t.tv_nsec * 1e-6;
}

static uint32_t opt_u32(const char *env, uint32_t dflt)
{
    const char *v = getenv(env);
    return (v && *v) ? (uint32_t)strtoul(v, NULL, 10) : dflt;
}

static int cmp_double(const void *a, const void *b)
{
    double x = *(const double *)a, y = *(const double *)b;
    return (x > y) - (x < y);
}

static void *xmalloc(size_t n)
{
    void *p = malloc(n ? n : 1);
    if (!p) {
        fprintf(stderr, "emb_host: out of memory (%zu bytes)\n", n);
        exit(12);
    }
    return p;
}

int main(int argc, char **argv)
{
    uint32_t T = opt_u32("NR_TABLES", 9), D = opt_u32("NR_COLS", 64);
    uint32_t B = opt_u32("MAX_NR_BATCHES", 64), P = opt_u32("MAX_INDICES_PER_BATCH", 32);
    uint64_t R = 50000;
    int iters = 100, validate = 1;
    for (int i = 1; i < argc; i++) {
        const char *a = argv[i];
        const char *v = (i + 1 < argc) ? argv[i + 1] : NULL;
        if (!strcmp(a, "--no-validate")) {
            validate = 0;
            continue;
        }
        if (!v) {
            fprintf(stderr, "emb_host: %s needs a value\n", a);
            return 2;
        }
        if (!strcmp(a, "--tables")) T = (uint32_t)strtoul(v, NULL, 10);
        else if (!strcmp(a, "--cols")) D = (uint32_t)strtoul(v, NULL, 10);
        else if (!strcmp(a, "--rows")) R = strtoull(v, NULL, 10);
        else if (!strcmp(a, "--batches")) B = (uint32_t)strtoul(v, NULL, 10);
        else if (!strcmp(a, "--pooling")) P = (uint32_t)strtoul(v, NULL, 10);
        else if (!strcmp(a, "--iters")) iters = atoi(v);
        else if (!strcmp(a, "--seed")) rng_state ^= strtoull(v, NULL, 10);
        else {
            fprintf(stderr, "emb_host: unknown option %s\n", a);
            return 2;
        }
        i++;
    }
    load_library(argv[0]);
    if (!T || !D || !B || !P || !R || iters < 1) {
        fprintf(stderr, "emb_host: tables, cols, rows, batches, pooling, iters must be > 0\n");
        return 2;
    }
    int rc = emb_configure(T, D, B, P, -1);
    if (rc) {
        fprintf(stderr, "emb_host: emb_configure: %s\n", emb_last_error());
        return 1;
    }

    /* tables: random int32, kept row-major on the host for the check; pushed column by column
     * through populate_mram() as dlrm_dpu_pytorch.py:328-332 does */
    int32_t **tables = xmalloc(T * sizeof *tables);
    int32_t *column = xmalloc(R * sizeof *column);
    struct dpu_set_t *handle = NULL;
    dpu_runtime_totals rt;
    memset(&rt, 0, sizeof rt);
    for (uint32_t t = 0; t < T; t++) {
        tables[t] = xmalloc(R * D * sizeof(int32_t));
        for (uint64_t i = 0; i < R * D; i++)
            tables[t][i] = (int32_t)(uint32_t)rng_next();
        for (uint32_t j = 0; j < D; j++) {
            for (uint64_t r = 0; r < R; r++)
                column[r] = tables[t][r * D + j];
            handle = populate_mram(t, R, j, column, &rt);
        }
    }
    free(column);

    const uint64_t ilen = (uint64_t)B * P;
    uint32_t **indices = xmalloc(T * sizeof *indices), **offsets = xmalloc(T * sizeof *offsets);
    float **results = xmalloc(T * sizeof *results);
    for (uint32_t t = 0; t < T; t++) {
        indices[t] = xmalloc(ilen * 4);
        offsets[t] = xmalloc((size_t)B * 4);
        results[t] = xmalloc((size_t)B * D * 4);
        for (uint32_t b = 0; b < B; b++)
            offsets[t][b] = b * P;
        for (uint64_t i = 0; i < ilen; i++)
            indices[t][i] = (uint32_t)(rng_next() % R);
    }

    double *lat = xmalloc((size_t)iters * sizeof *lat);
    emb_timings tm;
    double k_ms = 0;
    lookup(indices, offsets, results, handle, 0); /* warm-up: first-call allocations */
    for (int i = 0; i < iters; i++) {
        double t0 = now_ms();
        lookup(indices, offsets, results, handle, 0);
        lat[i] = now_ms() - t0;
        emb_last_timings(&tm);
        k_ms += tm.kernel_ms;
    }
    qsort(lat, (size_t)iters, sizeof *lat, cmp_double);
    double median = lat[iters / 2];

    long bad = 0;
    if (validate) {
        for (uint32_t t = 0; t < T; t++)
            for (uint32_t b = 0; b < B; b++)
                for (uint32_t j = 0; j < D; j++) {
                    uint32_t s = 0;
                    for (uint32_t p = 0; p < P; p++)
                        s += (uint32_t)tables[t][(uint64_t)indices[t][(uint64_t)b * P + p] * D + j];
                    float want = (float)((double)(float)(int32_t)s / pow(10, 9));
                    if (memcmp(&want, &results[t][(uint64_t)b * D + j], 4) != 0)
                        bad++;
                }
    }
    double lookups = (double)T * B * P;
    printf("{\"bench\": \"emb_host\", \"tables\": %u, \"cols\": %u, \"rows\": %llu, \"batches\": %u, "
           "\"pooling\": %u, \"iters\": %d, \"lookup_ms_median\": %.4f, \"lookup_ms_min\": %.4f, "
           "\"kernel_ms_mean\": %.4f, \"lookups_per_s\": %.4e, \"populate_s\": %.3f, "
           "\"validated\": %s, \"mismatches\": %ld}\n",
           T, D, (unsigned long long)R, B, P, iters, median, lat[0], k_ms / iters,
           lookups / (median * 1e-3), rt.execution_time_populate_copy_in, validate ? "true" : "false", bad);
    emb_shutdown();
    return bad ? 1 : 0;
}
