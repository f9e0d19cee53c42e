/*
 * emb_b200.h -- C ABI of the B200 embedding-lookup backend (shared object: emblib.so).
 *
 * Drop-in boundary for the ONE hot path of UBC-ECE-Sasha/PIM-Embedding-Lookup: the
 * sum-pooled EmbeddingBag lookup behind populate_mram()/lookup().  Plain C: pointers and
 * sizes only, no CUDA or torch types (a cudaStream_t travels as void*).
 *
 * Part 1 keeps the reference's two exported entry points and its ABI-visible structs, bit
 * for bit.  Part 2 is new surface (explicit lengths, device-resident buffers, float
 * tables, row-range shards, peer output slabs) that the reference's callers do not know
 * about; nothing in part 1 depends on it.
 *
 * Reference paths are relative to the upstream repository root.
 */
#ifndef EMB_B200_H
#define EMB_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ======================================================================================
 * Part 1 -- the reference's plugin surface
 * ==================================================================================== */

/* upmem/include/emb_host.h:41-48; ctypes mirror upmem/dputypes.py:66-79.  Seconds. */
typedef struct dpu_runtime_totals {
    double execution_time_prepare;          /* host-side staging                    */
    double execution_time_populate_copy_in; /* populate_mram: host -> HBM           */
    double execution_time_copy_in;          /* lookup: indices/offsets host -> HBM   */
    double execution_time_copy_out;         /* lookup: pooled rows HBM -> host       */
    double execution_time_aggregate_result; /* dequantise (fused: always 0)          */
    double execution_time_launch;           /* kernel time                           */
} dpu_runtime_totals;

/* emb_host.h:54-57; dputypes.py:18-33 (note the field order: nsec first). */
typedef struct dpu_timespec {
    long tv_nsec;
    long tv_sec;
} dpu_timespec;

/* emb_host.h:63-66; dputypes.py:35-46 */
typedef struct dpu_runtime_interval {
    dpu_timespec start;
    dpu_timespec stop;
} dpu_runtime_interval;

/* emb_host.h:72-75; dputypes.py:12-16 */
typedef enum dpu_runtime_config { RT_ALL = 0, RT_LAUNCH = 1 } dpu_runtime_config;

/* emb_host.h:81-85; dputypes.py:48-64 */
typedef struct dpu_runtime_group {
    unsigned int in_use;
    unsigned int length;
    dpu_runtime_interval *intervals;
} dpu_runtime_group;

/* upmem/include/emb_types.h:21-24 (packed pair pushed to every DPU before a launch). */
struct query_len {
    uint32_t indices_len;
    uint32_t nr_batches;
} __attribute__((packed));

/* Opaque here: the reference hands the UPMEM SDK's set to Python as an integer and back
 * (dlrm_dpu_pytorch.py:327,332,564).  The B200 library returns the address of its
 * process-wide context under the same type name. */
struct dpu_set_t;

/*
 * Replaces upmem/include/emb_host.h:136-183.
 * Store column `col` of table `table_id` (nr_rows int32 fixed-point values, scale 1e9) in
 * HBM.  Called once per (table, column) by PIM-dlrm-new/dlrm_dpu_pytorch.py:328-332.
 * The first call initialises the device and reads the shape from the environment
 * variables NR_TABLES, NR_COLS, MAX_NR_BATCHES, MAX_INDICES_PER_BATCH (compile-time
 * macros in the reference, upmem/Makefile:69-81) unless emb_configure() ran before.
 * `runtime` may be NULL; when given, execution_time_populate_copy_in is accumulated.
 * table_data is caller-owned and may be freed on return.  Errors print and exit(1), as
 * DPU_ASSERT does.  Returns the context handle to pass to lookup().
 */
struct dpu_set_t *populate_mram(uint32_t table_id, uint64_t nr_rows, uint32_t col,
                                int32_t *table_data, dpu_runtime_totals *runtime);

/*
 * Replaces upmem/include/emb_host.h:234-404 (+ post_process :186-222 and the DPU program
 * upmem/src/dpu/emb_dpu_lookup.c:35-138).
 * For every table t < NR_TABLES: indices[t] holds MAX_NR_BATCHES*MAX_INDICES_PER_BATCH
 * uint32 row ids, offsets[t] holds MAX_NR_BATCHES bag start positions (bag 0 starts at 0
 * whatever offsets[t][0] says; the last bag ends at INDICES_LEN), final_results[t]
 * receives MAX_NR_BATCHES x NR_COLS floats, row-major:
 *     final_results[t][b*NR_COLS + j] = (float)(int32 wrap-sum of column j) / 1e9 .
 * All pointers are HOST pointers; the call is synchronous.  dpu_set_ptr is the value
 * populate_mram() returned (ignored beyond a sanity check).  latency_print == 1 prints the
 * reference's six stage-latency lines (emb_host.h:395-402).  Returns NULL like the
 * reference (`return 0`).
 */
int32_t *lookup(uint32_t **indices, uint32_t **offsets, float **final_results,
                void *dpu_set_ptr, int64_t latency_print);

/* ======================================================================================
 * Part 2 -- extensions (new names; all return 0 on success, a negative EMB_E* otherwise)
 * ==================================================================================== */

#define EMB_OK 0
#define EMB_EINVAL (-1)  /* bad argument / shape                    */
#define EMB_ENOMEM (-2)  /* host or device allocation failed        */
#define EMB_ECUDA (-3)   /* a CUDA call failed, see emb_last_error  */
#define EMB_ESTATE (-4)  /* not configured / table not populated    */
#define EMB_ERANGE (-5)  /* an index fell outside its table         */

#define EMB_MAX_PEERS 8

/* table element kinds */
#define EMB_TABLE_FIXED 1 /* int32 fixed point, scale 1e9 (the DPU representation)        */
#define EMB_TABLE_FLOAT 2 /* fp32 (the CPU path, nn.EmbeddingBag(mode="sum"))             */

/* emb_dev_query.flags */
#define EMB_Q_RAW_SUMS 1u /* FIXED tables: write the int32 sums, skip the /1e9 (row shards) */
/* Plans only, and only when EVERY query of the plan carries it: nothing this launch READS (tables, indices,
 * offsets) is written by the kernel in front of it on the stream.  The launch may then gather and pool while that
 * kernel is still running (programmatic dependent launch); only its result stores wait for it to complete.  In a
 * loop of emb_plan_launch [+ emb_exchange_finish] calls this hides the pooling launch's tail, the launch latency and
 * the exchange's completion wait behind the next batch's gathers.  Without the flag a launch reads nothing before
 * the kernel in front of it has completed (stream order as usual). */
#define EMB_Q_OVERLAP 2u
/* The indices of this query are already slab row ids of a table reordered with emb_table_remap() (the caller
 * translated them, e.g. with emb_remap_indices() in its data pipeline): the launch does not translate them again. */
#define EMB_Q_REMAPPED 4u

typedef struct emb_shape {
    uint32_t nr_tables;             /* NR_TABLES                                      */
    uint32_t nr_cols;               /* NR_COLS (embedding dimension)                   */
    uint32_t max_nr_batches;        /* MAX_NR_BATCHES (bags per table per lookup())   */
    uint32_t max_indices_per_batch; /* MAX_INDICES_PER_BATCH (pooling factor)         */
} emb_shape;

/* Runtime replacement for the reference's -D shape macros (upmem/Makefile:69-81).
 * device < 0: use $EMB_DEVICE, else $LOCAL_RANK, else 0.  May be called again to change
 * max_nr_batches / max_indices_per_batch; nr_cols cannot change once a table exists. */
int emb_configure(uint32_t nr_tables, uint32_t nr_cols, uint32_t max_nr_batches,
                  uint32_t max_indices_per_batch, int device);
int emb_get_shape(emb_shape *out);
int emb_device(void);

/* Several GPUs of one box behind the reference's two entry points (the reference spreads its tables over all DPU
 * ranks of the machine the same way; emb_host.h:136-183 allocates NR_TABLES*NR_COLS DPUs).  A DEVICE GROUP places
 * whole tables on its GPUs -- table t on GPU table_device[t] of the group, or t % n_devices when table_device is NULL
 * -- and then
 *   - populate_mram / emb_populate_table_* / emb_table_* / emb_free_table act on the GPU that holds the table;
 *   - lookup / emb_lookup_ex / emb_lookup_ex_async cut a call by GPU, enqueue every part, then wait: each GPU moves
 *     its tables' indices in and pooled rows out over its own PCIe link at the same time as the others.  The
 *     reference's per-table result arrays need no exchange, so there is no collective;
 *   - emb_check_indices / emb_shutdown cover every GPU; emb_last_timings reports the slowest GPU's stage times and
 *     the summed bytes and launches.
 * Device-resident entry points (plans, emb_lookup_device, the exchange, emb_device_alloc ...) act on the group's FIRST
 * GPU only (a call that takes a device pointer or a caller's stream for a table on another GPU of the group returns
 * EMB_EINVAL and says so); several GPUs with device-resident data are one process per GPU (ShardedTables,
 * INTEGRATION.md section 6).  emb_plan_launch / emb_finish_launch take no lock and use the GPU the plan / finish object
 * was made on, so they may run in one thread while another thread is inside a group's host-pointer call.
 * Every entry point leaves the calling thread's current CUDA device as it found it, whatever GPU it worked on.
 * An unmodified caller gets a group from the environment: EMB_DEVICES="0,1,2,3" or EMB_NGPUS=4, read where the shape
 * variables are read.  The group can change only after emb_shutdown().  devices[] may name a GPU more than once (two
 * contexts on one GPU: of use for tests only). */
int emb_configure_devices(uint32_t nr_tables, uint32_t nr_cols, uint32_t max_nr_batches,
                          uint32_t max_indices_per_batch, const int *devices, uint32_t n_devices,
                          const uint32_t *table_device);
int emb_group_size(void);                /* GPUs in the group (1 by default)                  */
int emb_table_device(uint32_t table_id); /* CUDA device that holds (or will hold) the table   */
/* Frees every table, buffer, plan-independent resource; the next call re-initialises.
 * (The reference never frees its DPU set, emb_host.h:156-160.) */
void emb_shutdown(void);
const char *emb_last_error(void);

/* ---- tables ------------------------------------------------------------------------ */

/* Whole-table load from HOST memory, row-major [nr_rows][nr_cols].
 * emb_populate_table_i32: already-quantised int32 (what nr_cols populate_mram calls build).
 * emb_populate_table_f32 with kind EMB_TABLE_FIXED runs the reference's export arithmetic
 *   (dlrm_dpu_pytorch.py:318-320: float32 product by 1e9, rint, int64, low 32 bits) on the
 *   GPU; with EMB_TABLE_FLOAT the fp32 values are kept as they are. */
int emb_populate_table_i32(uint32_t table_id, uint64_t nr_rows, const int32_t *row_major);
int emb_populate_table_f32(uint32_t table_id, uint64_t nr_rows, const float *row_major, int kind);
/* Same, source already in DEVICE memory (copied; `stream` is a cudaStream_t or NULL). */
int emb_populate_table_device(uint32_t table_id, uint64_t nr_rows, const void *dev_row_major,
                              int src_is_f32, int kind, void *stream);
/* Row-range shard: the table slot holds global rows [row_base, row_base + nr_rows).
 * Indices outside the range contribute nothing (and are not errors). */
int emb_table_set_row_base(uint32_t table_id, uint64_t row_base);
/* Copy rows [row0, row0+n) back to the host as stored (int32 or fp32), row-major. */
int emb_table_read(uint32_t table_id, uint64_t row0, uint64_t n, void *dst_host);
int emb_table_info(uint32_t table_id, uint64_t *nr_rows, int *kind, const void **dev_ptr);
int emb_free_table(uint32_t table_id);
/* Mark a byte range of a table as L2-persisting for subsequent launches on `stream`
 * (cudaAccessPolicyWindow); nbytes == 0 clears the window. */
int emb_l2_persist(uint32_t table_id, uint64_t first_row, uint64_t nr_rows, float hit_ratio,
                   void *stream);

/* ---- hot rows: skewed index distributions --------------------------------------------- */

/* north_star: "an L2 persisting-access window keeps hot rows resident under skewed index distributions".  A
 * cudaAccessPolicyWindow (emb_l2_persist above) is ONE address range per stream, so it cannot serve a launch that
 * pools many tables.  These calls give every table its own hot set instead: the pooling kernel loads slab rows
 * [0, hot_rows) of a table with one L2 eviction policy and the rest with another (createpolicy + ld.global.nc
 * .L2::cache_hint, per load), and emb_table_remap() moves the rows a caller knows to be hot into that prefix.
 * Results are identical with and without (parity: tests/test_gpu_hot.py); what changes is which rows L2 keeps.
 * Measurements: profiles/r2_l2_hot_rows.md. */
#define EMB_HOT_KEEP 1        /* rows [0, hot_rows): L2 evict_last                         */
#define EMB_HOT_STREAM_COLD 2 /* rows [hot_rows, nr_rows): L2 evict_first                  */
/* mode = 0 or a combination of the two; hot_rows is clamped to the table.  Takes effect for plans created and
 * host-pointer / one-shot lookups issued afterwards; cleared when the table is populated again. */
int emb_table_set_hot_rows(uint32_t table_id, uint64_t hot_rows, int mode);
/* Reorder a populated (whole, not row-sharded) table in HBM: the row the caller calls r moves to slab row
 * new_of_old[r] (HOST array of nr_rows entries, a permutation of 0..nr_rows-1; may be freed on return).  Callers
 * keep using THEIR row ids: lookup(), emb_lookup_ex() and plans translate every index through the permutation
 * while they stage it (4 more bytes read per lookup), unless a query says EMB_Q_REMAPPED.  Needs a second slab of
 * the table's size while it runs; plans made before it refuse to launch (EMB_ESTATE).  Populating the table again
 * restores the caller's order.  emb_table_read() returns rows in slab order. */
int emb_table_remap(uint32_t table_id, const uint32_t *new_of_old);
/* HOST helper (no device needed): the permutation that puts the most looked-up rows first.  counts[r] = how often
 * row r was looked up in a sample of the workload; equal counts keep their order. */
int emb_hot_order(const uint64_t *counts, uint64_t nr_rows, uint32_t *new_of_old);
/* Translate n row ids in DEVICE memory to slab order, in place, on `stream` (for EMB_Q_REMAPPED queries).  A table
 * that was not reordered leaves them unchanged. */
int emb_remap_indices(uint32_t table_id, uint32_t *indices_dev, uint64_t n, void *stream);

/* ---- lookups with explicit lengths, HOST pointers ------------------------------------ */

/* Like lookup() but per-table lengths are arguments (the historical signatures in
 * upmem/c_test.py:48 and dlrm_dpu_pytorch.py:631-640 carried them too), so pooling may
 * vary per bag and per call.  Tables 0..nr_tables-1.  Bag b of table t covers
 * indices[t][start_b .. end_b) with start_0 = 0, start_b = offsets[t][b],
 * end_b = offsets[t][b+1], end_last = indices_len[t]. */
int emb_lookup_ex(uint32_t nr_tables, const uint32_t *const *indices,
                  const uint32_t *const *offsets, const uint32_t *indices_len,
                  const uint32_t *nr_batches, float *const *final_results);

/* The same call without the wait: everything is enqueued (copies in, pooling launches, copies out -- or the mapped
 * / bounced launch of a small call) and the function returns a ticket; emb_lookup_wait(ticket) returns once the
 * pooled rows are in final_results[].  TWO calls may be in flight: issued as
 *     async(k+1); wait(k); <consume k>; async(k+2); wait(k+1); ...
 * the copy-in of call k+1 crosses PCIe while call k is being pooled and its rows drain, so both PCIe directions stay
 * busy across calls, not only across the table groups inside one call (C2 on page-locked arrays: one call per
 * ~2.2 ms against 2.75 ms for back-to-back synchronous calls; profiles/r2_host_async.md).  Rules: all host arrays
 * of a call (pointer tables excluded: they are read before the function returns) stay untouched until its wait
 * returns; two calls in flight use different result arrays; issuing a third call first completes the oldest one
 * (its later wait returns at once); lookup() / emb_lookup_ex() / emb_shutdown() complete whatever is in flight
 * first.  emb_last_timings() describes the call completed last.  Thread-safe; a wait blocks outside the
 * library's lock. */
typedef uint64_t emb_ticket;
int emb_lookup_ex_async(uint32_t nr_tables, const uint32_t *const *indices,
                        const uint32_t *const *offsets, const uint32_t *indices_len,
                        const uint32_t *nr_batches, float *const *final_results, emb_ticket *ticket);
int emb_lookup_wait(emb_ticket ticket);

/* How lookup() / emb_lookup_ex() move the caller's HOST arrays across PCIe:
 * EMB_HOST_COPY        all staged: cudaMemcpyAsync of indices/offsets into HBM, pooling kernel,
 *                      cudaMemcpyAsync of the pooled rows back, pipelined over groups of
 *                      tables.  Works for any host memory (pageable included).
 * EMB_HOST_MAPPED_OUT  inputs staged as above, but the pooling kernel stores the pooled rows
 *                      straight into final_results[t] over PCIe (no device->host copies).
 *                      Needs page-locked final_results arrays.
 * EMB_HOST_MAPPED      nothing staged: the kernel also reads indices/offsets in place.  One
 *                      launch per call; the lowest latency for small calls.  Needs every array
 *                      page-locked.
 * EMB_HOST_BOUNCE      any host memory: the CPU copies the arrays into one page-locked arena, the
 *                      MAPPED launch runs on the arena, the CPU copies the results out.  For
 *                      small calls (the reference's own shapes) this is one CUDA launch instead
 *                      of three copies per table.
 * EMB_HOST_AUTO        (default) BOUNCE for tiny calls (<= 64 KiB) whatever the arrays are; MAPPED when
 *                      every array is page-locked and the inputs are small (<= 1 MiB); else BOUNCE when
 *                      the whole call is <= max(1 MiB, 128 KiB x tables); else COPY, cut into ~32 MB groups
 *                      of tables (one group for an asynchronous call issued under a call in flight).
 *                      Large calls are copied both ways also when the result arrays are page-locked:
 *                      measured faster than MAPPED_OUT (profiles/r2_host_path.md).
 * "Page-locked" = cudaHostAlloc, cudaHostRegister, emb_host_alloc or emb_host_register memory.
 * A forced path whose requirement is not met returns EMB_EINVAL. */
#define EMB_HOST_AUTO 0
#define EMB_HOST_COPY 1
#define EMB_HOST_MAPPED 2
#define EMB_HOST_MAPPED_OUT 3
#define EMB_HOST_BOUNCE 4
int emb_set_host_path(int path);

/* ---- device-resident lookups (plans) -------------------------------------------------- */

typedef struct emb_dev_query {
    uint32_t table_id;
    uint32_t nr_batches;      /* bags                                                  */
    uint32_t indices_len;     /* end of the last bag                                   */
    uint32_t flags;           /* EMB_Q_*                                               */
    const uint32_t *indices;  /* DEVICE, uint32 [indices_len]                           */
    const uint32_t *offsets;  /* DEVICE, uint32 [nr_batches]                            */
    void *out;                /* DEVICE, bag b -> out + b*out_stride elements (4 B each) */
    uint32_t out_stride;      /* elements between bags; 0 means nr_cols                 */
    /* When bags_per_rank != 0 the output of bag b goes to the slab of rank b / bags_per_rank:
     * out_peer[b / bags_per_rank] + (b % bags_per_rank)*out_stride.  The slabs may be
     * peer-GPU memory mapped with emb_ipc_open(): the pooled rows then cross NVLink straight
     * from the pooling kernel (fused all-to-all). */
    uint32_t bags_per_rank;
    void *out_peer[EMB_MAX_PEERS];
    /* Bag slice: pool only bags [bag_first, bag_first + bag_count) of the nr_batches bags the arrays
     * describe (bag_count == 0: all from bag_first on).  Bag numbering, bag bounds and output rows
     * stay those of the whole batch, so several ranks holding a replica of a small table can each
     * take a slice of the batch (data-parallel placement, no partial sums). */
    uint32_t bag_first;
    uint32_t bag_count;
    /* Order only, never results: pool the bags starting from bag_rotate (a bag id inside the slice,
     * rounded down to the launch's bags-per-CTA) and wrap around.  With bags_per_rank set, rank r
     * passes r * bags_per_rank so that at any moment every destination slab is being written by ONE
     * sender instead of all of them (all ranks otherwise walk the destinations in the same order and
     * N-1 senders pile onto one GPU's NVLink ingress). */
    uint32_t bag_rotate;
    uint32_t reserved_;
} emb_dev_query;

typedef struct emb_plan emb_plan;

/* Validate + upload the descriptors once; launches are then a single kernel with no copies. */
int emb_plan_create(const emb_dev_query *queries, uint32_t n, emb_plan **out);
int emb_plan_launch(emb_plan *plan, void *stream);
void emb_plan_destroy(emb_plan *plan);
/* Algorithmic bytes of one launch: L*(4D+4) + bags*(4D+4); needs host-known
 * sizes only. */
uint64_t emb_plan_bytes(const emb_plan *plan);
uint32_t emb_plan_grid(const emb_plan *plan);
/* One-shot: upload the descriptors and launch on `stream`, asynchronously (no synchronise, no
 * allocation); ordered with the caller's other work on that stream.  NULL stream = the library's
 * own stream (emb_stream_sync(NULL) waits for it). */
int emb_lookup_device(const emb_dev_query *queries, uint32_t n, void *stream);
/* Non-zero once any launch has seen an out-of-range index (cleared by the call; syncs). */
int emb_check_indices(void);

/* ---- a lookup that allocates nothing --------------------------------------------------------- */

/* No counterpart in the reference (its host library allocates DPUs and transfer buffers as it pleases).
 *
 * Who this is for.  Every other lookup here leans on memory the library got from cudaMalloc: the table slabs, the
 * descriptor buffers the pooling kernel reads its work list from, the word an out-of-range index sets.  Some callers
 * cannot afford that:
 *   - a plugin in a process whose own allocator owns the GPU (a serving engine or a framework caching allocator that
 *     reserves almost all of HBM at start-up): a cudaMalloc from a library then fails, or takes memory the host
 *     application had counted on;
 *   - a latency-bound serving loop: cudaMalloc / cudaFree stall the whole device, and the other entry points
 *     re-allocate whenever a call outgrows the previous one.
 * emb_lookup_static() never calls cudaMalloc, cudaFree, cudaHostAlloc, nor creates streams or events.  EVERYTHING it
 * touches is the caller's -- tables included, so populate_mram() / emb_configure() are not needed (and not used) --
 * except the kernel's work list and error word, which sit in a fixed array compiled into emblib.so (static
 * __device__, 11.5 KB).  Every GPU the library is loaded on carries that array until the process exits.
 *
 * The price of a fixed array:
 *   - at most EMB_STATIC_MAX_QUERIES queries (tables) per call; more is refused with EMB_EINVAL and a message that
 *     says so -- cut the call in two;
 *   - all calls on a GPU share the one array, and it keeps what the previous call left in it.  A call rewrites the
 *     descriptors it uses and clears the error word itself, on `stream`, before its pooling launch, so calls issued
 *     one after the other on ONE stream (or otherwise ordered on the GPU) are fine.  Two calls that may execute at
 *     the same time -- different streams without an event between them, or two threads -- overwrite each other's
 *     descriptors: wrong results or wild reads.  The library orders the enqueues of concurrent threads but cannot
 *     order the GPU; that is the caller's job;
 *   - the first call on each GPU looks up the array's device address (cudaGetSymbolAddress; the call works on the
 *     calling thread's CURRENT device, and the arrays and `stream` must belong to it).  Make that first call during
 *     start-up, outside any stream capture; later calls on that GPU only launch kernels.
 *
 * Arithmetic and bag bounds are those of lookup() / emb_lookup_ex(); the kernel is the index-ordered one (tuning
 * variant 1), so fp32 tables are summed strictly in index order.  An index outside [0, nr_rows) adds nothing. */
#define EMB_STATIC_MAX_QUERIES 64

typedef struct emb_static_query {
    const void *table;       /* DEVICE, row-major [nr_rows][nr_cols], 4-byte elements; 16-byte aligned when nr_cols
                                is 4, 8, 16, 32, 64 or 128                                              */
    uint64_t nr_rows;        /* 1 .. 2^32-2                                                             */
    uint32_t kind;           /* EMB_TABLE_FIXED or EMB_TABLE_FLOAT                                      */
    uint32_t flags;          /* 0 or EMB_Q_RAW_SUMS                                                     */
    uint32_t nr_batches;     /* bags                                                                    */
    uint32_t indices_len;    /* end of the last bag                                                     */
    const uint32_t *indices; /* DEVICE, uint32 [indices_len]                                            */
    const uint32_t *offsets; /* DEVICE, uint32 [nr_batches]                                             */
    void *out;               /* DEVICE, bag b -> out + b*out_stride elements (4 B each)                 */
    uint32_t out_stride;     /* elements between bags; 0 means nr_cols                                  */
    uint32_t reserved_;
} emb_static_query;

/* Enqueue on `stream` (a cudaStream_t; NULL = the legacy default stream -- this entry point owns no stream).
 * out_of_range == NULL: returns as soon as the kernels are enqueued.  Otherwise the call waits for `stream` and stores
 * 1 there if an index of THIS call fell outside its table, else 0. */
int emb_lookup_static(const emb_static_query *queries, uint32_t n, uint32_t nr_cols, void *stream,
                      int *out_of_range);

/* out[b*out_stride + j] = combine over s < n_parts of parts[s][b*part_stride + j]:
 * raw int32 partials are wrap-added then divided by 1e9 (is_raw_i32 = 1); fp32 partials are
 * added in order s = 0,1,...  Used at the receiving rank for row-split tables. */
int emb_reduce_partials(const void *const *parts_dev, uint32_t n_parts, uint32_t part_stride,
                        void *out_dev, uint32_t out_stride, uint64_t nr_bags, int is_raw_i32,
                        void *stream);

/* ---- completion step of the fused exchange (one launch per batch) --------------------------- */

/* One row-split table at the receiving rank: out[b*out_stride + j] = combine over s < n_parts of
 * parts[s][b*part_stride + j], b < nr_bags, j < nr_cols (same arithmetic as emb_reduce_partials). */
typedef struct emb_reduce_job {
    const void *parts[EMB_MAX_PEERS]; /* DEVICE                                            */
    void *out;                        /* DEVICE                                            */
    uint32_t n_parts;
    uint32_t part_stride; /* elements; 0 means nr_cols                                     */
    uint32_t out_stride;  /* elements; 0 means nr_cols                                     */
    uint32_t is_raw_i32;
    uint64_t nr_bags;
} emb_reduce_job;

typedef struct emb_finish emb_finish;

/* flag_blocks[r] = rank r's flag block (EMB_MAX_PEERS uint32, zeroed before first use, in memory
 * from emb_device_alloc) as mapped in this process (emb_ipc_open for r != my_rank).  jobs may be
 * NULL when n_jobs == 0 (no row-split table). */
int emb_finish_create(void *const *flag_blocks, uint32_t n_ranks, uint32_t my_rank,
                      const emb_reduce_job *jobs, uint32_t n_jobs, emb_finish **out);
/* ONE kernel on `stream`, after the pooling launch of the same batch: signal every peer's flag
 * block with `epoch`, wait until every peer has signalled `epoch` here, then run the reduce jobs.
 * `epoch` must be the same on every rank for the same batch and grow by one per batch (start at
 * 1).  A peer that has not arrived after 10 s sets the error word (emb_check_indices() != 0)
 * instead of hanging the GPU. */
int emb_finish_launch(emb_finish *f, uint32_t epoch, void *stream);
void emb_finish_destroy(emb_finish *f);

/* ---- memory helpers -------------------------------------------------------------------- */

int emb_device_alloc(void **dev_ptr, size_t nbytes); /* cudaMalloc (IPC-exportable)         */
int emb_device_free(void *dev_ptr);
int emb_host_alloc(void **host_ptr, size_t nbytes);  /* pinned                              */
int emb_host_free(void *host_ptr);
/* Page-lock memory the caller already owns (cudaHostRegister, mapped + portable) so that
 * lookup()/emb_lookup_ex() can write results into it / read inputs from it in place.  The range
 * must stay allocated until emb_host_unregister(). */
int emb_host_register(void *host_ptr, size_t nbytes);
int emb_host_unregister(void *host_ptr);
int emb_memcpy_h2d(void *dst_dev, const void *src_host, size_t nbytes, void *stream);
int emb_memcpy_d2h(void *dst_host, const void *src_dev, size_t nbytes, void *stream);
int emb_stream_sync(void *stream);

#define EMB_IPC_HANDLE_BYTES 64
typedef struct emb_ipc_handle {
    unsigned char bytes[EMB_IPC_HANDLE_BYTES];
} emb_ipc_handle;
/* Export a buffer from emb_device_alloc() / open a peer's export in this process. */
int emb_ipc_export(void *dev_ptr, emb_ipc_handle *out);
int emb_ipc_open(const emb_ipc_handle *h, void **peer_dev_ptr);
int emb_ipc_close(void *peer_dev_ptr);

/* ---- timings ----------------------------------------------------------------------------- */

typedef struct emb_timings {
    double h2d_ms;    /* indices + offsets host -> device   */
    double kernel_ms; /* pooling kernel                     */
    double d2h_ms;    /* pooled rows device -> host         */
    double total_ms;  /* wall clock of the whole call       */
    uint64_t h2d_bytes;
    uint64_t d2h_bytes;
    uint64_t launches; /* kernels launched by the call       */
    uint64_t path;     /* EMB_HOST_COPY / _MAPPED / _MAPPED_OUT / _BOUNCE: how the arrays crossed PCIe */
} emb_timings;
/* Stage times of the last lookup()/emb_lookup_ex() call (CUDA events). */
int emb_last_timings(emb_timings *out);
/* Kernels this library has launched since load (bench.py's gpu_launches). */
uint64_t emb_launch_count(void);
/* Tuning knob, mostly for measurement: rows in flight per sub-warp (4 or 8); 0 = default. */
int emb_set_unroll(int unroll);
/* All tuning knobs at once (0 = default): variant 1 = one bag per sub-warp (fp32 sums strictly in index
 * order; also used when a launch mixes fp32 and int32 tables), 4 = TMA gather4 (rows of 64..512 bytes),
 * 7 = the short-bag kernel (a warp per 32 bags, no shared memory; what the default picks by itself for launches that
 * average at most 1.25 indices per bag -- the reference's Kaggle shape, one index per bag; fp32 in index order),
 * 6 = the stretch kernel (default: load-balanced, no atomics; int32 bit-exact, fp32 deterministic and within
 * 1e-5 relative of the sequential sum); unroll = rows in flight per lane (4 or 8); bags_mult = bags per CTA
 * in units of sub-warps per CTA (1, 2, 4 or 8; variant 6).  Variants 2, 3 and 5 were round-1 kernels and no
 * longer exist.  Takes effect for plans created and host-pointer lookups issued afterwards. */
int emb_set_tuning(int variant, int unroll, int bags_mult);

#ifdef __cplusplus
}
#endif
#endif /* EMB_B200_H */
