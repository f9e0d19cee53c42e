/*
 * emb_oracle.c -- CPU restatement of the reference's embedding-lookup arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the shipped product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this file's shared object, and there only as the checker (or as the
 * timed CPU baseline), never as the thing measured or shipped.
 *
 * Parity status: PINNED.  Every function below is checked (tests/test_oracle_golden.py)
 * against the .npz files under tests/golden/, which oracle/make_golden.py produced by running the
 * reference's own sources -- upmem/src/dpu/emb_dpu_lookup.c and upmem/include/emb_host.h,
 * compiled verbatim from /root/reference against the UPMEM-SDK stand-in in
 * oracle/upmem_sim/ -- plus the reference's toy known-answer vector (upmem/c_test.py:40,56).
 *
 * Each function cites the reference lines it restates (paths relative to /root/reference).
 * "Intended semantics" = SURVEY.md Appendix A: the reference behaviour on the shapes where
 * it is well defined (even B <= 512, NR_COLS == 64, T*D <= 256), extended unchanged to
 * every shape.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include <pthread.h>
#include <stdlib.h>
#include <unistd.h>

/* ---- tiny pthread parallel-for (the image's default gcc has no libgomp) ------------ */

static int g_threads = 0; /* 0 = all online cores */

typedef void (*emb_range_fn)(int64_t lo, int64_t hi, void *ctx);

struct emb_job {
    emb_range_fn fn;
    void *ctx;
    int64_t lo, hi;
};

static void *emb_job_main(void *p)
{
    struct emb_job *j = (struct emb_job *)p;
    j->fn(j->lo, j->hi, j->ctx);
    return NULL;
}

int emb_oracle_num_threads(void)
{
    if (g_threads > 0)
        return g_threads;
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}

void emb_oracle_set_num_threads(int n) { g_threads = n > 0 ? n : 0; }

/* static split of [0, n) over the configured threads; the caller's thread takes chunk 0 */
static void emb_parallel_for(int64_t n, emb_range_fn fn, void *ctx)
{
    int nt = emb_oracle_num_threads();
    if (nt > 256)
        nt = 256;
    if (n < 2 * (int64_t)nt || nt <= 1) {
        fn(0, n, ctx);
        return;
    }
    pthread_t tid[256];
    struct emb_job job[256];
    int started[256];
    for (int i = 0; i < nt; i++) {
        job[i].fn = fn;
        job[i].ctx = ctx;
        job[i].lo = n * i / nt;
        job[i].hi = n * (i + 1) / nt;
        started[i] = 0;
    }
    for (int i = 1; i < nt; i++)
        started[i] = pthread_create(&tid[i], NULL, emb_job_main, &job[i]) == 0;
    fn(job[0].lo, job[0].hi, ctx);
    for (int i = 1; i < nt; i++) {
        if (started[i])
            pthread_join(tid[i], NULL);
        else
            fn(job[i].lo, job[i].hi, ctx); /* could not spawn: run it here */
    }
}

/* ---- a1: table quantisation ------------------------------------------------------- */

/* PIM-dlrm-new/dlrm_dpu_pytorch.py:318
 *     tmp_emb = np.rint(W.numpy() * (10**9)).astype(int)
 * W is float32 and the Python int is weakly typed, so the product is a *float32*
 * multiply; np.rint rounds half-to-even in float32; astype(int) is int64; the later
 * (c_int32 * n)(*col) (:329) keeps the low 32 bits. */
int32_t emb_oracle_quantise_one(float w)
{
    volatile float prod = w * 1.0e9f;      /* volatile: forbid fused/extended evaluation */
    float r = rintf(prod);                 /* FE_TONEAREST == half-to-even == np.rint */
    int64_t wide = (int64_t)r;
    return (int32_t)(uint32_t)(uint64_t)wide;
}

/* dlrm_dpu_pytorch.py:318-320: quantise, then transpose to [D][R] contiguous columns
 * (np.transpose + np.ascontiguousarray), each column being one populate_mram() payload. */
void emb_oracle_quantise_columns(const float *w_rowmajor, uint64_t nr_rows, uint32_t nr_cols,
                                 int32_t *cols_out /* [nr_cols][nr_rows] */)
{
    for (uint64_t r = 0; r < nr_rows; r++)
        for (uint32_t c = 0; c < nr_cols; c++)
            cols_out[(uint64_t)c * nr_rows + r] = emb_oracle_quantise_one(w_rowmajor[r * nr_cols + c]);
}

/* Same values, kept row-major [R][D] (the layout load_generator.c:40-65 validates with). */
void emb_oracle_quantise_rows(const float *w_rowmajor, uint64_t n_elems, int32_t *q_out)
{
    for (uint64_t i = 0; i < n_elems; i++)
        q_out[i] = emb_oracle_quantise_one(w_rowmajor[i]);
}

/* ---- a7: the DPU pooling loop, one (table, column) -------------------------------- */

/* upmem/src/dpu/emb_dpu_lookup.c:60-63,106-116,129: bag i sums the column entries named
 * by indices[start_i .. end_i) with int32 wrap-around, where
 *   start_0 = 0 (offsets[0] is never read, :60-63), start_i = offsets[i] (:60-61,:126-127)
 *   end_i   = offsets[i+1] for i < nr_batches-1, indices_len for the last bag (:109-110).
 * A bag whose end is not above its start contributes 0 (the while condition is false). */
void emb_oracle_pool_column(const int32_t *column, const uint32_t *indices,
                            const uint32_t *offsets, uint32_t indices_len, uint32_t nr_batches,
                            int32_t *results /* [nr_batches] */)
{
    for (uint32_t i = 0; i < nr_batches; i++) {
        uint32_t p = (i == 0) ? 0u : offsets[i];
        uint32_t end = (i == nr_batches - 1) ? indices_len : offsets[i + 1];
        uint32_t acc = 0; /* unsigned: defined wrap-around == DPU int32 add */
        for (; p < end; p++)
            acc += (uint32_t)column[indices[p]];
        results[i] = (int32_t)acc;
    }
}

/* ---- a5: post_process dequantisation ---------------------------------------------- */

/* upmem/include/emb_host.h:210
 *     final_results[t][k*NR_COLS+j] = (float)tmp_results[t][j][k] / pow(10,9);
 * int32 -> float (RNE), promoted to double for the divide by pow(10,9) == 1e9 exactly,
 * result rounded to float by the store. */
float emb_oracle_dequant(int32_t s)
{
    return (float)((double)(float)s / pow(10, 9));
}

/* ---- a4: a whole lookup() call, column-major tables -------------------------------- */

/* emb_host.h:234-404 + emb_dpu_lookup.c + post_process, intended semantics:
 * cols[t] points at D columns of R_t int32 each (the populate_mram payloads, stored
 * [D][R_t]); for every table the same (indices, offsets) pair goes to its D column DPUs
 * (emb_host.h:258-270); every DPU pools with emb_oracle_pool_column; post_process writes
 * final_results[t][k*D + j] (emb_host.h:207-212).  indices_len/nr_batches are per call
 * (INDICES_LEN / MAX_NR_BATCHES in the reference, emb_host.h:282-283). */
void emb_oracle_lookup_columns(uint32_t nr_tables, uint32_t nr_cols, const uint64_t *nr_rows,
                               const int32_t *const *cols, const uint32_t *const *indices,
                               const uint32_t *const *offsets, const uint32_t *indices_len,
                               const uint32_t *nr_batches, float *const *final_results)
{
    for (uint32_t t = 0; t < nr_tables; t++) {
        uint32_t nb = nr_batches[t];
        for (uint32_t j = 0; j < nr_cols; j++) {
            const int32_t *column = cols[t] + (uint64_t)j * nr_rows[t];
            for (uint32_t k = 0; k < nb; k++) {
                uint32_t p = (k == 0) ? 0u : offsets[t][k];
                uint32_t end = (k == nb - 1) ? indices_len[t] : offsets[t][k + 1];
                uint32_t acc = 0;
                for (; p < end; p++)
                    acc += (uint32_t)column[indices[t][p]];
                final_results[t][(uint64_t)k * nr_cols + j] = emb_oracle_dequant((int32_t)acc);
            }
        }
    }
}

/* ---- second, independent statement: row-major tables ------------------------------- */

/* upmem/src/load_generator.c:40-65 (validate_result): recompute the per-bag int32 sums
 * from row-major tables emb_tables[i][index*nr_cols + t].  Restated with the bag bounds of
 * the DPU kernel (validate_result reads offsets[i][j+1] one past the end for the last
 * bag; the intended bound there is indices_len, :51).  Also returns the raw int32 sums
 * when sums_out != NULL.  Bags are split statically over pthreads (emb_oracle_set_num_threads). */
struct rows_ctx {
    uint32_t nr_cols;
    const int32_t *tab;
    const uint32_t *idx, *off;
    uint32_t indices_len, nb;
    float *out;
    int32_t *sums;
};

static void rows_range(int64_t lo, int64_t hi, void *p)
{
    const struct rows_ctx *c = (const struct rows_ctx *)p;
    uint32_t nr_cols = c->nr_cols;
    uint32_t acc[1024]; /* nr_cols > 1024 never happens on this path (NR_COLS is 16..128) */
    for (int64_t k = lo; k < hi; k++) {
        uint32_t q = (k == 0) ? 0u : c->off[k];
        uint32_t end = (k == (int64_t)c->nb - 1) ? c->indices_len : c->off[k + 1];
        memset(acc, 0, sizeof(uint32_t) * nr_cols);
        for (; q < end; q++) {
            const int32_t *row = c->tab + (uint64_t)c->idx[q] * nr_cols;
            for (uint32_t j = 0; j < nr_cols; j++)
                acc[j] += (uint32_t)row[j];
        }
        for (uint32_t j = 0; j < nr_cols; j++) {
            if (c->sums)
                c->sums[(uint64_t)k * nr_cols + j] = (int32_t)acc[j];
            c->out[(uint64_t)k * nr_cols + j] = emb_oracle_dequant((int32_t)acc[j]);
        }
    }
}

void emb_oracle_lookup_rows(uint32_t nr_tables, uint32_t nr_cols, const int32_t *const *tables,
                            const uint32_t *const *indices, const uint32_t *const *offsets,
                            const uint32_t *indices_len, const uint32_t *nr_batches,
                            float *const *final_results, int32_t *const *sums_out)
{
    if (nr_cols > 1024)
        abort();
    for (uint32_t t = 0; t < nr_tables; t++) {
        struct rows_ctx c = { nr_cols, tables[t], indices[t], offsets[t], indices_len[t],
                              nr_batches[t], final_results[t], sums_out ? sums_out[t] : NULL };
        emb_parallel_for(nr_batches[t], rows_range, &c);
    }
}

/* ---- a9: the float path (nn.EmbeddingBag(mode="sum") on CPU) ----------------------- */

/* PIM-dlrm-new/dlrm_s_pytorch.py:295,483-488 -> ATen index_select_add ->
 * PIM-Pytorch/caffe2/perfkernels/embedding_lookup_idx.cc:33-77 (generic kernel):
 *   out[b][:] = 0;  for p in bag (index order):  out[b][j] += 1.0f * W[idx_p][j]
 * fp32 adds, sequential in index order, no reassociation.  Same bag bounds as above
 * (torch requires offsets[0] == 0 and ends the last bag at len(indices)). */
struct f32_ctx {
    const float *table;
    uint32_t nr_cols;
    const uint32_t *idx, *off;
    uint32_t indices_len, nb;
    float *out;
};

static void f32_range(int64_t lo, int64_t hi, void *p)
{
    const struct f32_ctx *c = (const struct f32_ctx *)p;
    for (int64_t k = lo; k < hi; k++) {
        uint32_t q = (k == 0) ? 0u : c->off[k];
        uint32_t end = (k == (int64_t)c->nb - 1) ? c->indices_len : c->off[k + 1];
        float *o = c->out + (uint64_t)k * c->nr_cols;
        for (uint32_t j = 0; j < c->nr_cols; j++)
            o[j] = 0.0f;
        for (; q < end; q++) {
            const float *row = c->table + (uint64_t)c->idx[q] * c->nr_cols;
            for (uint32_t j = 0; j < c->nr_cols; j++)
                o[j] = o[j] + row[j];
        }
    }
}

void emb_oracle_bag_sum_f32(const float *table, uint32_t nr_cols, const uint32_t *indices,
                            const uint32_t *offsets, uint32_t indices_len, uint32_t nr_batches,
                            float *out /* [nr_batches][nr_cols] */)
{
    struct f32_ctx c = { table, nr_cols, indices, offsets, indices_len, nr_batches, out };
    emb_parallel_for(nr_batches, f32_range, &c);
}
