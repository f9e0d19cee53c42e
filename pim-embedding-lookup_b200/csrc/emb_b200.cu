/*
 * emb_b200.cu -- host side of emblib.so: the C ABI declared in include/emb_b200.h.
 *
 * Replaces the body of upmem/include/emb_host.h (populate_mram :136-183, lookup :234-404,
 * post_process :186-222).  Where the reference allocates DPUs, pushes per-DPU buffers over
 * the UPMEM rank transfer and launches tasklets, this file keeps tables resident in HBM,
 * moves a call's indices/offsets with cudaMemcpyAsync and launches one pooling kernel
 * (emb_kernels.cuh) for all tables.  There is no CPU fallback: without a CUDA device every
 * entry point fails loudly.
 */
#include "emb_kernels.cuh"

#include <cuda.h>  // CUtensorMap types only: cuTensorMapEncodeTiled is fetched with cudaGetDriverEntryPoint

#include <errno.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <utility>
#include <vector>

namespace {

using emb::BagDesc;

constexpr uint32_t kMaxTables = 1u << 16;
constexpr int kRingSlots = 4;

struct Table {
    void *dev = nullptr;
    uint64_t rows = 0;
    uint64_t row_base = 0;
    int kind = 0;  // 0 = empty, EMB_TABLE_FIXED, EMB_TABLE_FLOAT
    bool sharded = false;
    void *d_tmap = nullptr;  // CUtensorMap [rows][D] in device memory, built on first use by the TMA kernel
    uint32_t generation = 0; // bumped whenever the slab is (re)allocated or freed: plans remember it
    // hot rows (emb_table_set_hot_rows / emb_table_remap): slab rows [0, hot_rows) carry the L2 policy hot_mode names;
    // d_remap != NULL: the slab was reordered, caller row r lives at slab row d_remap[r]
    uint32_t *d_remap = nullptr;
    uint64_t hot_rows = 0;
    uint32_t hot_mode = 0;
};

/* What host_complete() needs to know about a host-pointer call whose work has been enqueued. */
struct HostPending {
    bool active = false;
    uint64_t ticket = 0;  // 0: a synchronous call
    double t_begin = 0;
    uint32_t T = 0, n_groups = 0;
    bool timed_in = false;  // the inputs went through the copy-in stream (its events were recorded)
    bool in_mapped = false, out_mapped = false, bounce = false, desc_same = false;
    bool use_done_event = false;  // wait for HostSlot::ev_done, not for the whole stream (another call may follow on it)
    cudaStream_t s_last = nullptr;
    uint64_t h2d_payload = 0, d2h_payload = 0, launches = 0;
    size_t desc_bytes = 0;
    // bounced calls: the pooled rows of table t lie at h_bounce + b_src[t] and go to b_dst[t]
    std::vector<float *> b_dst;
    std::vector<size_t> b_src, b_bytes;
};

/* Everything one host-pointer call owns while it is in flight: device mirrors of the caller's arrays, descriptors,
 * events, the bounce arena.  The synchronous entry points use slot 0; emb_lookup_ex_async alternates between the
 * slots so that the copies and kernels of two consecutive calls overlap. */
struct HostSlot {
    uint32_t *d_idx = nullptr;
    uint32_t *d_off = nullptr;
    float *d_out = nullptr;
    size_t idx_cap = 0, off_cap = 0, out_cap = 0;  // elements
    BagDesc *d_descs = nullptr;
    BagDesc *h_descs = nullptr;  // pinned
    uint32_t desc_cap = 0;
    // bytes d_descs holds (last successful upload): a caller that passes the same arrays call after
    // call (dlrm_dpu_pytorch.py:484-498 allocates them once) needs no descriptor upload at all
    std::vector<uint8_t> desc_shadow;
    bool desc_shadow_ok = false;
    std::vector<cudaEvent_t> ev_pool;
    cudaEvent_t ev_done = nullptr;
    // page-locked bounce arena for small calls on pageable arrays
    uint8_t *h_bounce = nullptr;
    uint8_t *d_bounce = nullptr;  // device alias of h_bounce
    size_t bounce_cap = 0;
    HostPending p;
};
constexpr int kHostSlots = 2;

struct Ctx {
    bool configured = false;
    bool device_ready = false;
    int device = 0;
    int sm_count = 148;
    emb_shape shape = {0, 0, 0, 0};
    std::vector<Table> tables;

    cudaStream_t stream = nullptr;     // compute + H2D
    cudaStream_t stream_out = nullptr; // D2H
    cudaStream_t stream_in = nullptr;  // H2D of the pipelined host-pointer path
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    uint32_t *d_err = nullptr;

    // populate_mram() column staging
    void *d_col = nullptr;
    size_t col_bytes = 0;

    // host-pointer lookup path: per-call state, two slots (emb_lookup_ex_async keeps two calls in flight)
    HostSlot hslot[kHostSlots];
    // descriptor ring of emb_lookup_device
    BagDesc *ring_h[4] = {nullptr, nullptr, nullptr, nullptr};
    BagDesc *ring_d[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ring_ev[4] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t ring_cap = 0;
    int ring_next = 0;
    emb_timings last = {0, 0, 0, 0, 0, 0, 0, 0};
};

/* Process-wide knobs (emb_set_tuning, emb_set_host_path); 0 = default.  They survive emb_shutdown(). */
struct Tuning {
    int unroll = 0;     // rows in flight per sub-warp: 4 or 8
    int variant = 0;    // 1 = bag-per-sub-warp kernel, 4 = TMA gather4, 6 = stretch kernel
    int bags_mult = 0;  // stretch kernel: bags per CTA = mult * sub-warps per CTA (1, 2, 4, 8)
    int host_path = EMB_HOST_AUTO;  // emb_set_host_path
};
Tuning g_tune;

/* One context per GPU of the device group (emb_configure_devices / $EMB_NGPUS; a group of one by default).  Every
 * function below works on "the current context" `g`; entry points that name a table switch to the context of the GPU
 * that holds it (UseCtx), the host-pointer lookups cut a call by GPU and run the parts concurrently.  Everything is
 * serialised by g_mu, so the current context is a plain global. */
constexpr int kMaxDevices = EMB_MAX_PEERS;
Ctx g_ctxs[kMaxDevices];
Ctx *g_cur = &g_ctxs[0];
#define g (*g_cur)
struct Group {
    int n = 1;                      // contexts in use: g_ctxs[0 .. n)
    std::vector<int> table_device;  // explicit placement (index into the group), empty = table_id % n
};
Group g_grp;
uint64_t g_next_ticket = 1;  // tickets of emb_lookup_ex_async: slot = ticket % kHostSlots, on every GPU of the group
emb_timings g_group_last = {0, 0, 0, 0, 0, 0, 0, 0};  // emb_last_timings() of a call that ran on several GPUs
bool g_group_last_valid = false;
std::mutex g_mu;
std::atomic<uint64_t> g_launches{0};
thread_local char g_err[512] = "";

int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(EMB_ECUDA, "%s:%d: %s -> %s", __FILE__, __LINE__, #call,                   \
                        cudaGetErrorString(e_));                                                   \
    } while (0)

/* The reference aborts on any SDK error (DPU_ASSERT, emb_host.h:157-173); its two entry
 * points have no error return, so ours do the same. */
[[noreturn]] void die(const char *where, int code)
{
    fprintf(stderr, "emblib(B200): %s failed (%d): %s\n", where, code, g_err);
    exit(EXIT_FAILURE);
}

uint32_t env_u32(const char *name, uint32_t dflt)
{
    const char *v = getenv(name);
    if (!v || !*v)
        return dflt;
    char *end = nullptr;
    unsigned long x = strtoul(v, &end, 10);
    return (end && *end == 0) ? (uint32_t)x : dflt;
}

double now_s()
{
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return t.tv_sec + t.tv_nsec * 1e-9;
}

/* The calling thread's CUDA device, saved at the top of every entry point that takes the lock and put back when it
 * returns.  In between the thread works on the current context's GPU: a caller whose own current device is another
 * one (torch.cuda.set_device(1) beside EMB_DEVICE=0, a second thread at its default device 0 beside a library bound
 * to GPU 2, any thread beside a device group) would otherwise have the library's allocations, events and copies land
 * on the caller's device -- and would get its device switched behind its back by a group call.  Declared after the
 * lock_guard, so it is destroyed (device restored) before the lock is released. */
struct DeviceScope {
    int saved = -1;
    DeviceScope()
    {
        if (cudaGetDevice(&saved) != cudaSuccess) {
            saved = -1;
            cudaGetLastError();  // no device visible: the entry point reports that itself
        } else if (g_cur->device_ready && saved != g_cur->device) {
            cudaSetDevice(g_cur->device);
        }
    }
    ~DeviceScope()
    {
        int now = -1;
        if (saved >= 0 && cudaGetDevice(&now) == cudaSuccess && now != saved)
            cudaSetDevice(saved);
    }
    DeviceScope(const DeviceScope &) = delete;
    DeviceScope &operator=(const DeviceScope &) = delete;
};

/* For the entry points that take no lock (emb_plan_launch, emb_finish_launch): launch on the GPU the object was made
 * on even when the calling thread's current device is another one, and leave that device as it was. */
struct OnDevice {
    int saved = -1;
    explicit OnDevice(int device)
    {
        if (cudaGetDevice(&saved) != cudaSuccess || saved == device)
            saved = -1;
        else
            cudaSetDevice(device);
    }
    ~OnDevice()
    {
        if (saved >= 0)
            cudaSetDevice(saved);
    }
    OnDevice(const OnDevice &) = delete;
    OnDevice &operator=(const OnDevice &) = delete;
};

/* Make context k (and its GPU) the current one for the lifetime of the object.  Used inside a DeviceScope only, which
 * has put the thread on the previous context's GPU (and restores the caller's own device at the very end). */
struct UseCtx {
    Ctx *prev;
    explicit UseCtx(int k) : prev(g_cur)
    {
        g_cur = &g_ctxs[k];
        if (g_cur != prev && g_cur->device_ready)
            cudaSetDevice(g_cur->device);
    }
    ~UseCtx()
    {
        if (g_cur != prev && prev->device_ready)
            cudaSetDevice(prev->device);
        g_cur = prev;
    }
    UseCtx(const UseCtx &) = delete;
    UseCtx &operator=(const UseCtx &) = delete;
};

/* Which GPU of the group holds a table: whole tables, dealt round-robin unless the caller gave a map. */
int owner_of(uint32_t table_id)
{
    if (g_grp.n <= 1)
        return 0;
    if (table_id < g_grp.table_device.size())
        return g_grp.table_device[table_id];
    return (int)(table_id % (uint32_t)g_grp.n);
}

int ensure_device()
{
    if (g.device_ready)
        return EMB_OK;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0)
        return fail(EMB_ECUDA, "no CUDA device visible (%s): this library has no CPU path",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    if (g.device < 0 || g.device >= n)
        return fail(EMB_EINVAL, "device %d out of range (%d visible)", g.device, n);
    CU(cudaSetDevice(g.device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, g.device));
    g.sm_count = prop.multiProcessorCount;
    if (prop.major < 10)
        fprintf(stderr, "emblib(B200): warning: device %d is sm_%d%d; kernels are built for sm_100a\n",
                g.device, prop.major, prop.minor);
    CU(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&g.stream_out, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&g.stream_in, cudaStreamNonBlocking));
    for (auto &ev : g.ev)
        CU(cudaEventCreate(&ev));
    CU(cudaMalloc(&g.d_err, sizeof(uint32_t)));
    CU(cudaMemset(g.d_err, 0, sizeof(uint32_t)));
    g.device_ready = true;
    return EMB_OK;
}

int pick_device(int device)
{
    if (device >= 0)
        return device;
    const char *names[] = {"EMB_DEVICE", "LOCAL_RANK"};
    for (const char *nm : names) {
        const char *v = getenv(nm);
        if (v && *v)
            return atoi(v);
    }
    return 0;
}

int configure_locked(uint32_t T, uint32_t D, uint32_t B, uint32_t P, int device)
{
    if (T == 0 || T > kMaxTables || D == 0 || D > 4096)
        return fail(EMB_EINVAL, "bad shape: NR_TABLES=%u NR_COLS=%u", T, D);
    if ((uint64_t)B * P > 0xffffffffull)
        return fail(EMB_EINVAL, "MAX_NR_BATCHES*MAX_INDICES_PER_BATCH overflows uint32");
    if (g.configured && g.shape.nr_cols != D) {
        for (const Table &t : g.tables)
            if (t.kind)
                return fail(EMB_ESTATE, "NR_COLS cannot change from %u to %u while tables exist",
                            g.shape.nr_cols, D);
    }
    int dev = pick_device(device);
    if (g.device_ready && dev != g.device)
        return fail(EMB_ESTATE, "already bound to device %d", g.device);
    g.device = dev;
    g.shape = {T, D, B, P};
    if (g.tables.size() < T)
        g.tables.resize(T);
    g.configured = true;
    return ensure_device();
}

/* Configure a device group: context k on devices[k], all with the same shape.  A group can only change while no
 * context holds a device (emb_shutdown() first). */
int configure_group(uint32_t T, uint32_t D, uint32_t B, uint32_t P, const int *devices, int n, const uint32_t *table_device)
{
    if (!devices || n < 1 || n > kMaxDevices)
        return fail(EMB_EINVAL, "a device group has 1..%d GPUs", kMaxDevices);
    bool live = false;
    for (const Ctx &c : g_ctxs)
        live |= c.device_ready;
    if (live) {
        bool same = n == g_grp.n;
        for (int k = 0; k < n && same; k++)
            same = g_ctxs[k].device_ready && g_ctxs[k].device == devices[k];
        if (!same)
            return fail(EMB_ESTATE, "the device group cannot change while the library holds devices: emb_shutdown() first");
    }
    if (table_device)
        for (uint32_t t = 0; t < T; t++)
            if (table_device[t] >= (uint32_t)n)
                return fail(EMB_EINVAL, "table %u placed on GPU %u of a group of %d", t, table_device[t], n);
    for (int k = 0; k < n; k++) {
        if (devices[k] < 0)
            return fail(EMB_EINVAL, "device group member %d: device %d", k, devices[k]);
        UseCtx u(k);
        int rc = configure_locked(T, D, B, P, devices[k]);
        if (rc)
            return rc;
    }
    g_grp.n = n;
    g_grp.table_device.clear();
    if (table_device)
        g_grp.table_device.assign(table_device, table_device + T);
    if (g_cur->device_ready)
        cudaSetDevice(g_cur->device);
    return EMB_OK;
}

/* $EMB_DEVICES ("0,1,2,3") or $EMB_NGPUS (devices 0..n-1): the device group of a caller that knows nothing about
 * this library beyond populate_mram() / lookup().  Returns the number of devices (0: no group asked for). */
int group_from_env(int *devices)
{
    const char *v = getenv("EMB_DEVICES");
    int n = 0;
    if (v && *v) {
        while (*v && n < kMaxDevices) {
            char *end = nullptr;
            long d = strtol(v, &end, 10);
            if (end == v)
                break;
            devices[n++] = (int)d;
            v = (*end == ',') ? end + 1 : end;
        }
        return n;
    }
    n = (int)std::min<uint32_t>(env_u32("EMB_NGPUS", 0), kMaxDevices);
    for (int k = 0; k < n; k++)
        devices[k] = k;
    return n;
}

/* The reference bakes these four in with -D at build time (upmem/Makefile:69-81) and its
 * Python reads NR_TABLES from the environment (dlrm_dpu_pytorch.py:128); the run scripts
 * export all of them (upmem/run.sh:40-45,93-101).  Same names, read at first use. */
int ensure_configured()
{
    if (g.configured)
        return ensure_device();
    if (g_cur != &g_ctxs[0])
        return fail(EMB_ESTATE, "GPU %d of the device group is not configured", (int)(g_cur - g_ctxs));
    uint32_t T = env_u32("NR_TABLES", 0), D = env_u32("NR_COLS", 0);
    uint32_t B = env_u32("MAX_NR_BATCHES", 0), P = env_u32("MAX_INDICES_PER_BATCH", 0);
    if (!T || !D)
        return fail(EMB_ESTATE, "shape unknown: set NR_TABLES, NR_COLS, MAX_NR_BATCHES, "
                                "MAX_INDICES_PER_BATCH in the environment or call emb_configure()");
    int devices[kMaxDevices];
    const int n = group_from_env(devices);
    if (n > 1)
        return configure_group(T, D, B, P, devices, n, nullptr);
    return configure_locked(T, D, B, P, n == 1 ? devices[0] : -1);
}

int table_slot(uint32_t table_id, Table **out)
{
    if (table_id >= kMaxTables)
        return fail(EMB_EINVAL, "table_id %u too large", table_id);
    if (table_id >= g.tables.size())
        g.tables.resize(table_id + 1);
    *out = &g.tables[table_id];
    return EMB_OK;
}

/* (Re)allocate the slab of a table slot for `rows` rows.  The slab holds rows + 1 rows: the extra last row is
 * all zero and is what the stretch kernel gathers for an index that must contribute nothing (out of range, a
 * row of another shard, padding), so its row loads need no predicate.  Everything this function enqueues goes
 * on `s`, the stream the caller populates on, so the zero fill is ordered before the caller's copies.  A slot
 * that is reused forgets that it once was a row shard. */
int table_alloc(Table *t, uint64_t rows, int kind, cudaStream_t s)
{
    if (rows == 0 || rows > 0xfffffffeull)
        return fail(EMB_EINVAL, "nr_rows %llu unsupported (1 .. 2^32-2)", (unsigned long long)rows);
    t->row_base = 0;
    t->sharded = false;
    // rows arrive in the caller's order again: a reordering or a hot prefix of what the slot held before is void
    t->hot_rows = 0;
    t->hot_mode = 0;
    if (t->d_remap) {
        cudaDeviceSynchronize();  // plans translating through it may still be running
        cudaFree(t->d_remap);
        t->d_remap = nullptr;
        t->generation++;
    }
    if (t->dev && t->rows == rows) {
        t->kind = kind;
        return EMB_OK;
    }
    if (t->dev) {
        // plans made for the old slab may still be running on any stream
        cudaDeviceSynchronize();
        cudaFree(t->dev);
        t->dev = nullptr;
    }
    if (t->d_tmap) {
        cudaFree(t->d_tmap);
        t->d_tmap = nullptr;
    }
    t->generation++;
    size_t bytes = (size_t)(rows + 1) * g.shape.nr_cols * 4;
    cudaError_t e = cudaMalloc(&t->dev, bytes);
    if (e != cudaSuccess) {
        t->dev = nullptr;
        t->kind = 0;
        return fail(EMB_ENOMEM, "cudaMalloc(%zu bytes) for a table: %s", bytes, cudaGetErrorString(e));
    }
    // columns arrive one populate_mram() call at a time: unwritten ones read as 0; the last row stays 0
    e = cudaMemsetAsync(t->dev, 0, bytes, s);
    if (e != cudaSuccess)
        return fail(EMB_ECUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
    t->rows = rows;
    t->kind = kind;
    return EMB_OK;
}

int ensure_events(HostSlot &hs, size_t n)
{
    while (hs.ev_pool.size() < n) {
        cudaEvent_t e;
        CU(cudaEventCreate(&e));
        hs.ev_pool.push_back(e);
    }
    if (!hs.ev_done)
        CU(cudaEventCreateWithFlags(&hs.ev_done, cudaEventDisableTiming));
    return EMB_OK;
}

int grow(void **p, size_t *cap_elems, size_t want_elems, size_t elem)
{
    if (*cap_elems >= want_elems && *p)
        return EMB_OK;
    if (*p)
        cudaFree(*p);
    *p = nullptr;
    size_t n = std::max(want_elems, (size_t)256);
    cudaError_t e = cudaMalloc(p, n * elem);
    if (e != cudaSuccess)
        return fail(EMB_ENOMEM, "cudaMalloc(%zu bytes): %s", n * elem, cudaGetErrorString(e));
    *cap_elems = n;
    return EMB_OK;
}

/* 2-D tensor map [rows][D] (4-byte elements, box = one row) for the gather4 kernel.  Encoded on the
 * host by the driver (entry point fetched through the runtime: no link-time libcuda dependency)
 * and kept in device memory beside the table. */
int ensure_tmap(Table *t)
{
    if (t->d_tmap)
        return EMB_OK;
    typedef CUresult (*EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeTiled enc = nullptr;
    if (!enc) {
        cudaDriverEntryPointQueryResult qr;
        void *fn = nullptr;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
        if (e != cudaSuccess || !fn)
            return fail(EMB_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
        enc = (EncodeTiled)fn;
    }
    const uint32_t D = g.shape.nr_cols;
    if (t->rows > 0x7fffffffull)
        return fail(EMB_EINVAL, "the TMA kernel addresses rows with int32 coordinates: %llu rows is too many",
                    (unsigned long long)t->rows);
    CUtensorMap tm;
    memset(&tm, 0, sizeof tm);
    cuuint64_t gdim[2] = {D, t->rows};
    cuuint64_t gstr[1] = {(cuuint64_t)D * 4};
    cuuint32_t box[2] = {D, 1};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_INT32, 2, t->dev, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(EMB_ECUDA, "cuTensorMapEncodeTiled failed (%d) for a %llu x %u table", (int)r,
                    (unsigned long long)t->rows, D);
    CU(cudaMalloc(&t->d_tmap, sizeof tm));
    CU(cudaMemcpy(t->d_tmap, &tm, sizeof tm, cudaMemcpyHostToDevice));
    return EMB_OK;
}

uint32_t lanes_for_dim(uint32_t D)
{
    if (D % 4 != 0)
        return 0;
    uint32_t l = D / 4;
    if (l > 32 || (l & (l - 1)))
        return 0;
    return l;
}

/* Which kernel pools a launch, decided when the descriptors are built (the CTA -> bag mapping depends on it). */
struct Variant {
    int kind;      // 0 generic (any D), 1 bag-per-sub-warp (index-ordered adds, any table kind), 4 TMA gather4 (rows of
                   // 64..512 bytes, any table kind), 6 stretch kernel (load-balanced, no atomics; int32 or fp32 launches),
                   // 7 short-bag kernel (launches of about one index per bag; index-ordered adds, any table kind)
    int unroll;    // K: rows in flight per lane
    int mult;      // stretch kernel: bags per CTA / sub-warps per CTA
    uint32_t bags; // bags per CTA
    int stage;     // stretch kernel: 1 = the larger pass buffer (measurement knob)
    bool hot;      // stretch kernel: the HOT instantiation (hot-row tables among the launch's tables; K = 8)
};

/* Hot-row tables among the tables of a launch (any_hot: an L2 policy is set; any_remap: a reordered slab whose
 * indices the kernel must translate).  Only the stretch kernel knows them, as its HOT instantiation (K = 8, the small
 * pass); a launch that holds row shards keeps the compacting instantiation instead.  Policies are hints and may be
 * dropped; a translation may not. */
int apply_hot(Variant &v, bool any_hot, bool any_remap, bool compact)
{
    v.hot = false;
    if (!any_hot && !any_remap)
        return EMB_OK;
    if (v.kind == 6 && !compact) {
        v.hot = true;
        v.unroll = 8;
        v.stage = 0;
        return EMB_OK;
    }
    if (any_remap)
        return fail(EMB_EINVAL, "a reordered table (emb_table_remap) is pooled by the stretch kernel only (variant 6, "
                                "dimension 4..128 in powers of two, no row shards and no fp32/int32 mix in the launch)");
    return EMB_OK;
}

/* avg_pool = lookups per bag of the launch (0 = unknown).  The stretch kernel is the default; it replaced the
 * round-1 load-balanced kernels (per CTA with atomics, per warp) on every workload (profiles/r2_stretch_ab_rotated.txt).
 * Variant 1 stays for strictly index-ordered fp32 sums and for launches that mix fp32 and int32 tables, variant 4
 * (TMA gather4) as the measured alternative north_star asks for. */
int short_tpw()
{
    static const int tpw = [] {
        const uint32_t v = env_u32("EMB_SHORT_TPW", 2);  // measurement knob
        return v == 1 || v == 4 ? (int)v : 2;
    }();
    return tpw;
}

constexpr double kShortBagPool = 1.25;  // average indices per bag up to which a launch takes the short-bag kernel

Variant pick_variant(bool any_float, double avg_pool = 0.0, bool all_float = false, uint64_t total_bags = 0)
{
    Variant v;
    uint32_t D = g.shape.nr_cols;
    uint32_t lanes = lanes_for_dim(D);
    v.stage = 0;
    v.hot = false;
    if (!lanes) {
        v.kind = 0;
        v.unroll = 1;
        v.mult = 1;
        v.bags = emb::kThreads / 32;
        return v;
    }
    static const int dflt_kind = (int)env_u32("EMB_DEFAULT_VARIANT", 6);  // measurement knob
    v.kind = g_tune.variant ? g_tune.variant : (dflt_kind == 1 || dflt_kind == 4 ? dflt_kind : 6);
    // About one index per bag (the reference's Kaggle shape, README.md:14): the short-bag kernel, unless a table of this
    // GPU is reordered or carries an L2 policy (only the stretch kernel knows those).  EMB_SHORT_BAGS=0: measurement knob.
    static const bool short_on = env_u32("EMB_SHORT_BAGS", 1) != 0;
    if ((v.kind == 6 && !g_tune.variant && short_on && avg_pool > 0.0 && avg_pool <= kShortBagPool) || v.kind == 7) {
        bool special = false;
        for (const Table &t : g.tables)
            special |= t.d_remap != nullptr || (t.hot_mode != 0 && t.hot_rows != 0);
        if (!special) {
            v.kind = 7;
            v.unroll = (int)std::min<uint32_t>(lanes, 4);
            v.mult = 1;
            v.bags = emb::kThreads * (uint32_t)short_tpw();  // a warp owns 32 bags per tile whatever the dimension
            return v;
        }
        v.kind = 6;
    }
    if (v.kind == 4 && lanes < 4)
        v.kind = 6;  // the TMA kernel is for rows of 64 bytes and more
    if (any_float && v.kind != 4 && !(v.kind == 6 && all_float))
        v.kind = 1;  // fp32 sums in strict index order (embedding_lookup_idx.cc:33-77): kinds 1 and 4; the stretch
                     // kernel pools fp32 tables too (deterministic, reassociated at its cut points), but a launch
                     // is compiled for one element kind
    v.unroll = g_tune.unroll ? g_tune.unroll : (lanes >= 16 ? 8 : 4);
    v.mult = 1;
    v.bags = emb::kThreads / lanes;  // kind 1: one bag per sub-warp of a 256-thread CTA
    if (v.kind == 6) {
        // One warp per CTA.  Defaults from the round-2 sweeps on the B200 (tools/sweep.py,
        // profiles/r2_stretch_ab_rotated.txt; K = rows in flight per lane, mult = bags per CTA / sub-warps):
        //   64-byte rows (C2):   K8 m2 0.2257 ms | K4 m2 0.2360 | K8 m4 0.2545 (two passes) | K8 m1 0.2503
        //   ... one index per bag (C3): K4 m4 0.0353 | K8 m4 0.0373 | K8 m8 0.0414
        //   256-byte rows (C4):  K8 m4 0.5761 | K8 m2 0.5924 | K4 m4 0.6148 | K8 m8 0.6206
        //   512-byte rows (C5):  K8 m1 0.4222 | K8 m2 0.4469 | K8 m4 0.4828
        const bool short_bags = avg_pool > 0.0 && avg_pool <= 4.0;
        v.unroll = g_tune.unroll ? g_tune.unroll : (short_bags && lanes <= 8 ? 4 : 8);
        v.mult = g_tune.bags_mult ? g_tune.bags_mult : (lanes >= 32 ? 1 : lanes >= 16 ? 4 : short_bags ? 4 : 2);
        // Few bags (the reference's own shapes: RM2 as shipped is 32 tables x 64 bags x 120 lookups, upmem/r.sh:6-10):
        // the defaults above were swept at 65536 bags per table and would leave most SMs without a CTA -- 256 one-warp
        // CTAs walking 960 lookups each.  Fewer bags per CTA until the launch has a few CTAs per SM: a bag's lookups are
        // still spread over the sub-warps of its CTA, so long bags cost nothing extra.
        if (!g_tune.bags_mult && total_bags)
            while (v.mult > 1 && total_bags / (32 / lanes * (uint64_t)v.mult) < 8ull * (uint64_t)g.sm_count)
                v.mult /= 2;
        static const int big_stage = (int)env_u32("EMB_STRETCH_STAGE", 0);  // measurement knob: 1 = the larger pass
        v.stage = big_stage;
        v.bags = std::min<uint32_t>(32 / lanes * v.mult, 64);  // one marker byte per lookup names the bag
    }
    if (v.kind == 4)
        v.bags = (emb::kThreads / 32) * 4 * (32 / lanes);  // 4 bags per sub-warp
    return v;
}

template <int LANES>
void launch_ordered(const Ctx &c, int unroll, uint32_t grid, cudaStream_t s, const BagDesc *d, uint32_t n, uint32_t uni)
{
    if (unroll == 8)
        emb::bag_sum_kernel<LANES, 8><<<grid, emb::kThreads, 0, s>>>(d, n, uni, c.d_err);
    else
        emb::bag_sum_kernel<LANES, 4><<<grid, emb::kThreads, 0, s>>>(d, n, uni, c.d_err);
}

/* Stretch kernel: K rows in flight per lane, BAGS = sub-warps x mult bags per one-warp CTA, STAGE lookups per
 * pass (the smaller pass keeps shared memory -- and so the L1 the row loads live in -- small; knob
 * EMB_STRETCH_STAGE=1 doubles it). */
/* Launch with the programmatic-stream-serialization attribute: the kernel may become resident while the kernel in
 * front of it on the stream is draining (it must order itself with griddepcontrol.wait -- only kernels that do are
 * launched through here).  EMB_PDL=0 (measurement knob) launches plainly. */
template <typename... KArgs, typename... Args>
void launch_dependent(void (*kern)(KArgs...), dim3 grid, uint32_t block, cudaStream_t s, Args... args)
{
    static const bool pdl = env_u32("EMB_PDL", 1) != 0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(block, 1, 1);
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);  // the caller checks cudaGetLastError
}

constexpr int kHotMinBlocks = 32;  // resident one-warp CTAs per SM the HOT instantiation is compiled for

template <int LANES, int K, int MULT, bool IS_FLOAT>
void launch_stretch_m(const Ctx &c, int big_stage, uint32_t grid, cudaStream_t s, const BagDesc *d, uint32_t n, uint32_t uni, uint32_t overlap)
{
    constexpr int SUB = 32 / LANES;
    constexpr int BAGS = SUB * MULT < 64 ? SUB * MULT : 64;  // stretch_bags() below
    constexpr int UNIT = SUB * K;  // a pass is a whole number of K-lookup steps of every sub-warp
    constexpr int SMALL = (LANES <= 4 ? 256 : 128) < UNIT ? UNIT : (LANES <= 4 ? 256 : 128);
    constexpr int LARGE = 2 * SMALL;
    constexpr int MINB = 32;
    if (n > 65535)  // gridDim.y limit: search for the query instead
        uni = 0;
    const dim3 g2 = uni ? dim3(uni, n, 1) : dim3(grid, 1, 1);  // uniform launches: x = chunk, y = query
    // bit 1 of `overlap`: row shards among the descriptors -> the instantiation that stages only a shard's own
    // lookups (COMPACT; fixed-point tables, K = 8).  Anything else pools shards with the plain kernel, where a
    // foreign index gathers the all-zero row.
    if constexpr (!IS_FLOAT && K == 8) {
        if (overlap & 2u) {
            if (big_stage)
                launch_dependent(emb::bag_sum_stretch_kernel<LANES, K, BAGS, LARGE, IS_FLOAT, MINB, true>, g2, 32, s, d, n, uni,
                                 c.d_err, overlap & 1u);
            else
                launch_dependent(emb::bag_sum_stretch_kernel<LANES, K, BAGS, SMALL, IS_FLOAT, MINB, true>, g2, 32, s, d, n, uni,
                                 c.d_err, overlap & 1u);
            return;
        }
    }
    // bit 2: hot-row tables among the descriptors -> the HOT instantiation (K = 8, small pass; apply_hot)
    if constexpr (K == 8) {
        if (overlap & 4u) {
            launch_dependent(emb::bag_sum_stretch_kernel<LANES, K, BAGS, SMALL, IS_FLOAT, kHotMinBlocks, false, true>, g2, 32, s, d,
                             n, uni, c.d_err, overlap & 1u);
            return;
        }
    }
    overlap &= 1u;
    if (big_stage)
        launch_dependent(emb::bag_sum_stretch_kernel<LANES, K, BAGS, LARGE, IS_FLOAT, MINB>, g2, 32, s, d, n, uni, c.d_err,
                         overlap);
    else
        launch_dependent(emb::bag_sum_stretch_kernel<LANES, K, BAGS, SMALL, IS_FLOAT, MINB>, g2, 32, s, d, n, uni, c.d_err,
                         overlap);
}

template <int LANES, int K, bool IS_FLOAT>
void launch_stretch_k(const Ctx &c, const Variant &v, uint32_t grid, cudaStream_t s, const BagDesc *d, uint32_t n, uint32_t uni, uint32_t overlap)
{
    switch (v.mult) {
    case 1: return launch_stretch_m<LANES, K, 1, IS_FLOAT>(c, v.stage, grid, s, d, n, uni, overlap);
    case 4: return launch_stretch_m<LANES, K, 4, IS_FLOAT>(c, v.stage, grid, s, d, n, uni, overlap);
    case 8: return launch_stretch_m<LANES, K, 8, IS_FLOAT>(c, v.stage, grid, s, d, n, uni, overlap);
    default: return launch_stretch_m<LANES, K, 2, IS_FLOAT>(c, v.stage, grid, s, d, n, uni, overlap);
    }
}

template <int LANES>
void launch_stretch(const Ctx &c, const Variant &v, bool is_float, uint32_t grid, cudaStream_t s, const BagDesc *d, uint32_t n, uint32_t uni,
                    uint32_t overlap)
{
    if (is_float) {
        if (v.unroll == 4)
            return launch_stretch_k<LANES, 4, true>(c, v, grid, s, d, n, uni, overlap);
        return launch_stretch_k<LANES, 8, true>(c, v, grid, s, d, n, uni, overlap);
    }
    if (v.unroll == 4)
        return launch_stretch_k<LANES, 4, false>(c, v, grid, s, d, n, uni, overlap);
    launch_stretch_k<LANES, 8, false>(c, v, grid, s, d, n, uni, overlap);
}

template <int LANES>
void launch_tma(const Ctx &c, uint32_t grid, cudaStream_t s, const BagDesc *d, uint32_t n, uint32_t uni)
{
    if constexpr (LANES >= 4) {
        constexpr int WARPS = emb::kThreads / 32;
        constexpr size_t smem = (size_t)WARPS * emb::kTmaStages * emb::kTmaStageBytes + WARPS * emb::kTmaStages * 8 +
                                WARPS * (4 * (32 / LANES) + 1) * 4;
        static bool once[64] = {};  // a function attribute is per device
        if (c.device < 0 || c.device >= 64 || !once[c.device]) {
            cudaFuncSetAttribute(emb::bag_sum_tma_kernel<LANES, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (c.device >= 0 && c.device < 64)
                once[c.device] = true;
        }
        emb::bag_sum_tma_kernel<LANES, 3><<<grid, emb::kThreads, smem, s>>>(d, n, uni, c.d_err);
    } else {
        (void)grid; (void)s; (void)d; (void)n; (void)uni;
    }
}

template <int LANES>
void launch_short(const Ctx &c, uint32_t grid, cudaStream_t s, const BagDesc *d, uint32_t n, uint32_t uni)
{
    constexpr int UN = LANES < 4 ? LANES : 4;  // steps whose row loads are in flight together
    // Tiles of 32 bags per warp (short_tpw(), also the descriptors' bags per CTA), compiled for 4 resident CTAs per SM
    // (54-63 registers).  C3 on one box, isolated launch (profiles/r3_c3_short_bags.md): 1 tile 0.0366 ms, 2 tiles
    // 0.0314, 4 tiles 0.0325; compiled for 8 CTAs per SM (32 registers, spills) 0.0321 / 0.0321 / 0.0343; the stretch
    // kernel 0.0411.
    switch (short_tpw()) {
    case 1: emb::bag_sum_short_kernel<LANES, UN, 1, 4><<<grid, emb::kThreads, 0, s>>>(d, n, uni, c.d_err); break;
    case 4: emb::bag_sum_short_kernel<LANES, UN, 4, 4><<<grid, emb::kThreads, 0, s>>>(d, n, uni, c.d_err); break;
    default: emb::bag_sum_short_kernel<LANES, UN, 2, 4><<<grid, emb::kThreads, 0, s>>>(d, n, uni, c.d_err); break;
    }
}

template <int LANES>
void launch_for_lanes(const Ctx &c, const Variant &v, bool is_float, uint32_t grid, cudaStream_t s, const BagDesc *d, uint32_t n, uint32_t uni,
                      uint32_t overlap)
{
    if (v.kind == 7)
        launch_short<LANES>(c, grid, s, d, n, uni);
    else if (v.kind == 6)  // the one kernel that orders its result stores itself (EMB_Q_OVERLAP)
        launch_stretch<LANES>(c, v, is_float, grid, s, d, n, uni, overlap);
    else if (v.kind == 4)
        launch_tma<LANES>(c, grid, s, d, n, uni);
    else
        launch_ordered<LANES>(c, v.unroll, grid, s, d, n, uni);
}

/* Every query of the launch owns the same number of CTAs (the usual case: all tables pool the same
 * batch): CTA c belongs to descriptor c / chunks and no kernel has to search for it. */
uint32_t uniform_chunks(const BagDesc *h, uint32_t n)
{
    static const bool on = env_u32("EMB_UNIFORM_CHUNKS", 1) != 0;  // measurement knob
    if (!on || n == 0 || h[0].n_chunks == 0 || h[0].chunk_begin != 0)
        return 0;
    const uint32_t c = h[0].n_chunks;
    for (uint32_t i = 0; i < n; i++)
        if (h[i].n_chunks != c || h[i].chunk_begin != i * c)
            return 0;
    return c;
}

/* `uni` = uniform_chunks() of the n descriptors (0 = search). */
int launch_bag_sum(const Ctx &c, const Variant &v, const BagDesc *d_descs, uint32_t n, uint32_t uni, uint32_t grid, cudaStream_t s,
                   bool is_float = false, bool overlap = false, bool compact = false)
{
    if (grid == 0)
        return EMB_OK;
    uint32_t D = c.shape.nr_cols;
    const uint32_t obits = (overlap ? 1u : 0u) | (compact ? 2u : 0u) | (v.hot ? 4u : 0u);
    switch (v.kind ? lanes_for_dim(D) : 0) {
    case 0: emb::bag_sum_generic_kernel<<<grid, emb::kThreads, 0, s>>>(d_descs, n, uni, D, c.d_err); break;
    case 1: launch_for_lanes<1>(c, v, is_float, grid, s, d_descs, n, uni, obits); break;
    case 2: launch_for_lanes<2>(c, v, is_float, grid, s, d_descs, n, uni, obits); break;
    case 4: launch_for_lanes<4>(c, v, is_float, grid, s, d_descs, n, uni, obits); break;
    case 8: launch_for_lanes<8>(c, v, is_float, grid, s, d_descs, n, uni, obits); break;
    case 16: launch_for_lanes<16>(c, v, is_float, grid, s, d_descs, n, uni, obits); break;
    default: launch_for_lanes<32>(c, v, is_float, grid, s, d_descs, n, uni, obits); break;
    }
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());
    return EMB_OK;
}

/* Fill one descriptor from a query; returns the number of CTAs it needs. */
int fill_desc(const emb_dev_query &q, BagDesc *d, uint32_t chunk_begin, const Variant &variant, uint32_t *chunks)
{
    if (q.table_id >= g.tables.size() || !g.tables[q.table_id].kind)
        return fail(EMB_ESTATE, "table %u is not populated", q.table_id);
    Table &t = g.tables[q.table_id];
    const uint32_t per = variant.bags;
    uint32_t D = g.shape.nr_cols;
    memset(d, 0, sizeof *d);
    d->table = t.dev;
    if (variant.kind == 4) {
        int rc = ensure_tmap(&t);
        if (rc)
            return rc;
        d->tmap = t.d_tmap;
    }
    d->indices = q.indices;
    d->offsets = q.offsets;
    d->out = q.out;
    d->rows = (uint32_t)t.rows;
    d->row_base = (uint32_t)t.row_base;
    d->nr_batches = q.nr_batches;
    d->indices_len = q.indices_len;
    d->out_stride = q.out_stride ? q.out_stride : D;
    d->bags_per_rank = q.bags_per_rank;
    d->chunk_begin = chunk_begin;
    if (q.bag_first > q.nr_batches || q.bag_count > q.nr_batches - q.bag_first)
        return fail(EMB_EINVAL, "bag slice [%u, +%u) outside the %u bags of table %u", q.bag_first, q.bag_count,
                    q.nr_batches, q.table_id);
    d->bag_first = q.bag_first;
    d->bag_end = q.bag_count ? q.bag_first + q.bag_count : q.nr_batches;
    if (d->out_stride < D)
        return fail(EMB_EINVAL, "out_stride %u < NR_COLS %u", d->out_stride, D);
    uint32_t flags = 0;
    if (t.kind == EMB_TABLE_FLOAT)
        flags |= emb::kDescFloat;
    else if (q.flags & EMB_Q_RAW_SUMS)
        flags |= emb::kDescRaw;
    if (t.sharded)
        flags |= emb::kDescSharded;
    bool vec = d->out_stride % 4 == 0;
    if (q.nr_batches) {
        if (!q.offsets || (q.indices_len && !q.indices))
            return fail(EMB_EINVAL, "query for table %u has NULL indices/offsets", q.table_id);
        if (q.bags_per_rank) {
            uint32_t ranks = (q.nr_batches + q.bags_per_rank - 1) / q.bags_per_rank;
            if (ranks > EMB_MAX_PEERS)
                return fail(EMB_EINVAL, "nr_batches/bags_per_rank = %u ranks > %d", ranks, EMB_MAX_PEERS);
            for (uint32_t r = 0; r < ranks; r++) {
                if (!q.out_peer[r])
                    return fail(EMB_EINVAL, "out_peer[%u] is NULL", r);
                d->out_peer[r] = q.out_peer[r];
                vec = vec && ((uintptr_t)q.out_peer[r] % 16 == 0);
            }
        } else {
            if (!q.out)
                return fail(EMB_EINVAL, "query for table %u has NULL out", q.table_id);
            vec = vec && ((uintptr_t)q.out % 16 == 0);
        }
    }
    if (vec)
        flags |= emb::kDescVecStore;
    d->flags = flags;
    if (variant.hot) {
        d->remap = (q.flags & EMB_Q_REMAPPED) ? nullptr : t.d_remap;
        d->hot_mode = t.hot_mode;
        d->hot_rows = t.hot_mode ? (uint32_t)std::min<uint64_t>(t.hot_rows, t.rows) : 0;
    } else if (t.d_remap && !(q.flags & EMB_Q_REMAPPED)) {
        return fail(EMB_EINVAL, "table %u is reordered (emb_table_remap) but this launch cannot translate its indices",
                    q.table_id);
    }
    *chunks = (d->bag_end - d->bag_first + per - 1) / per;
    d->n_chunks = *chunks;
    // bag_rotate: start at the chunk holding that bag (rounded down to a chunk boundary) and wrap
    d->chunk_rot = 0;
    if (q.bag_rotate) {
        if (q.bag_rotate < d->bag_first || q.bag_rotate >= d->bag_end)
            return fail(EMB_EINVAL, "bag_rotate %u outside the query's bags [%u, %u)", q.bag_rotate, d->bag_first,
                        d->bag_end);
        d->chunk_rot = (q.bag_rotate - d->bag_first) / per;
    }
    return EMB_OK;
}

}  // namespace

struct emb_plan {
    std::vector<BagDesc> h;
    std::vector<std::pair<uint32_t, uint32_t>> gen;  // (table slot, Table::generation) of every query
    BagDesc *d = nullptr;
    // The context (GPU) the plan was made on -- the group's first GPU.  emb_plan_launch takes no lock and must not look at
    // the library's current context, which another thread's device-group call switches while it runs.
    Ctx *ctx = nullptr;
    uint32_t grid = 0;
    uint32_t uni = 0;  // uniform_chunks() of the descriptors
    uint64_t bytes = 0;
    Variant variant = {0, 1, 1, 8, 0, false};
    bool all_float = false;
    bool overlap = false;  // EMB_Q_OVERLAP on every query
    // The plan holds row shards of fixed-point tables: launched with the kernel instantiation that stages only a
    // shard's own lookups (bag_sum_stretch_kernel<..., COMPACT>; whole tables in the same launch are staged as usual).
    bool compact = false;
};

extern "C" {

static int host_drain();  // host-pointer calls still in flight complete first (defined with the host path below)

/* ---------------------------------------------------------------- configuration ------ */

int emb_configure(uint32_t nr_tables, uint32_t nr_cols, uint32_t max_nr_batches,
                  uint32_t max_indices_per_batch, int device)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    if (g_grp.n > 1) {  // a group exists: the shape changes on every GPU of it
        if (device >= 0 && device != g_ctxs[0].device)
            return fail(EMB_ESTATE, "a device group is configured: emb_shutdown() before binding to device %d", device);
        for (int k = 0; k < g_grp.n; k++) {
            UseCtx u(k);
            int rc = configure_locked(nr_tables, nr_cols, max_nr_batches, max_indices_per_batch, g.device);
            if (rc)
                return rc;
        }
        return EMB_OK;
    }
    if (!g.configured && device < 0) {
        int devices[kMaxDevices];
        const int n = group_from_env(devices);
        if (n > 1)
            return configure_group(nr_tables, nr_cols, max_nr_batches, max_indices_per_batch, devices, n, nullptr);
        if (n == 1)
            device = devices[0];
    }
    return configure_locked(nr_tables, nr_cols, max_nr_batches, max_indices_per_batch, device);
}

int emb_configure_devices(uint32_t nr_tables, uint32_t nr_cols, uint32_t max_nr_batches, uint32_t max_indices_per_batch,
                          const int *devices, uint32_t n_devices, const uint32_t *table_device)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    return configure_group(nr_tables, nr_cols, max_nr_batches, max_indices_per_batch, devices, (int)n_devices, table_device);
}

int emb_group_size(void) { return g_grp.n; }

int emb_table_device(uint32_t table_id)
{
    const int k = owner_of(table_id);
    return g_ctxs[k].device_ready ? g_ctxs[k].device : -1;
}

int emb_get_shape(emb_shape *out)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    if (!out)
        return fail(EMB_EINVAL, "NULL out");
    if (!g.configured)
        return fail(EMB_ESTATE, "not configured");
    *out = g.shape;
    return EMB_OK;
}

int emb_device(void) { return g.device_ready ? g.device : -1; }

const char *emb_last_error(void) { return g_err; }

uint64_t emb_launch_count(void) { return g_launches.load(); }

int emb_set_unroll(int unroll)
{
    if (unroll != 0 && unroll != 4 && unroll != 8)
        return fail(EMB_EINVAL, "unroll must be 0, 4 or 8");
    g_tune.unroll = unroll;
    return EMB_OK;
}

int emb_set_tuning(int variant, int unroll, int bags_mult)
{
    if (variant != 0 && variant != 1 && variant != 4 && variant != 6 && variant != 7)
        return fail(EMB_EINVAL, "variant must be 0 (auto), 1 (bag per sub-warp, index-ordered), 4 (TMA gather4), "
                                "6 (stretch kernel) or 7 (short-bag kernel); 2, 3 and 5 were round-1 kernels, removed");
    if (bags_mult != 0 && bags_mult != 1 && bags_mult != 2 && bags_mult != 4 && bags_mult != 8)
        return fail(EMB_EINVAL, "bags_mult must be 0, 1, 2, 4 or 8");
    int rc = emb_set_unroll(unroll);
    if (rc)
        return rc;
    g_tune.variant = variant;
    g_tune.bags_mult = bags_mult;
    return EMB_OK;
}

int emb_set_host_path(int path)
{
    if (path < EMB_HOST_AUTO || path > EMB_HOST_BOUNCE)
        return fail(EMB_EINVAL, "host path must be EMB_HOST_AUTO, _COPY, _MAPPED, _MAPPED_OUT or _BOUNCE");
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    g_tune.host_path = path;
    return EMB_OK;
}

static void shutdown_current()
{
    if (g.device_ready) {
        cudaSetDevice(g.device);
        host_drain();
        cudaDeviceSynchronize();
        for (Table &t : g.tables) {
            if (t.dev)
                cudaFree(t.dev);
            if (t.d_tmap)
                cudaFree(t.d_tmap);
            if (t.d_remap)
                cudaFree(t.d_remap);
        }
        cudaFree(g.d_col);
        cudaFree(g.d_err);
        for (HostSlot &hs : g.hslot) {  // a call still pending is abandoned: the device is idle by now
            cudaFree(hs.d_idx);
            cudaFree(hs.d_off);
            cudaFree(hs.d_out);
            cudaFree(hs.d_descs);
            if (hs.h_descs)
                cudaFreeHost(hs.h_descs);
            if (hs.h_bounce)
                cudaFreeHost(hs.h_bounce);
            for (auto &ev : hs.ev_pool)
                cudaEventDestroy(ev);
            if (hs.ev_done)
                cudaEventDestroy(hs.ev_done);
        }
        for (auto &ev : g.ev)
            if (ev)
                cudaEventDestroy(ev);
        for (int i = 0; i < kRingSlots; i++) {
            if (g.ring_h[i])
                cudaFreeHost(g.ring_h[i]);
            if (g.ring_d[i])
                cudaFree(g.ring_d[i]);
            if (g.ring_ev[i])
                cudaEventDestroy(g.ring_ev[i]);
        }
        if (g.stream)
            cudaStreamDestroy(g.stream);
        if (g.stream_out)
            cudaStreamDestroy(g.stream_out);
        if (g.stream_in)
            cudaStreamDestroy(g.stream_in);
    }
    g = Ctx();
}

void emb_shutdown(void)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    for (int k = kMaxDevices - 1; k >= 0; k--) {  // every GPU of the group; the tuning knobs survive
        g_cur = &g_ctxs[k];
        shutdown_current();
    }
    g_cur = &g_ctxs[0];
    g_grp = Group();
    g_group_last_valid = false;
}

/* ---------------------------------------------------------------- tables -------------- */

struct dpu_set_t *populate_mram(uint32_t table_id, uint64_t nr_rows, uint32_t col, int32_t *table_data,
                                dpu_runtime_totals *runtime)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    double t0 = now_s();
    int rc = ensure_configured();
    if (rc)
        die("populate_mram", rc);
    UseCtx on_owner(owner_of(table_id));  // the GPU of the device group that holds this table
    const uint32_t D = g.shape.nr_cols;
    if (col >= D || !table_data) {
        fail(EMB_EINVAL, "col %u >= NR_COLS %u or NULL data", col, D);
        die("populate_mram", EMB_EINVAL);
    }
    Table *t = nullptr;
    if ((rc = table_slot(table_id, &t)) || (rc = table_alloc(t, nr_rows, EMB_TABLE_FIXED, g.stream)))
        die("populate_mram", rc);
    size_t bytes = (size_t)nr_rows * 4;
    size_t cap = g.col_bytes;
    if ((rc = grow(&g.d_col, &cap, bytes, 1)))
        die("populate_mram", rc);
    g.col_bytes = cap;
    cudaError_t e = cudaMemcpyAsync(g.d_col, table_data, bytes, cudaMemcpyHostToDevice, g.stream);
    if (e == cudaSuccess) {
        uint32_t grid = (uint32_t)std::min<uint64_t>((nr_rows + 255) / 256, (uint64_t)g.sm_count * 16);
        emb::scatter_column_kernel<<<grid, 256, 0, g.stream>>>((uint32_t *)t->dev, (const uint32_t *)g.d_col,
                                                               nr_rows, D, col);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(g.stream);  // table_data may be freed on return
    if (e != cudaSuccess) {
        fail(EMB_ECUDA, "populate_mram: %s", cudaGetErrorString(e));
        die("populate_mram", EMB_ECUDA);
    }
    if (runtime)
        runtime->execution_time_populate_copy_in += now_s() - t0;
    return reinterpret_cast<struct dpu_set_t *>(&g_ctxs[0]);
}

static int populate_whole(uint32_t table_id, uint64_t nr_rows, const void *src, bool src_device,
                          bool src_f32, int kind, cudaStream_t s)
{
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!src)
        return fail(EMB_EINVAL, "NULL table data");
    if (kind != EMB_TABLE_FIXED && kind != EMB_TABLE_FLOAT)
        return fail(EMB_EINVAL, "kind must be EMB_TABLE_FIXED or EMB_TABLE_FLOAT");
    if (kind == EMB_TABLE_FLOAT && !src_f32)
        return fail(EMB_EINVAL, "a float table needs fp32 source data");
    Table *t = nullptr;
    if ((rc = table_slot(table_id, &t)) || (rc = table_alloc(t, nr_rows, kind, s)))
        return rc;
    uint64_t n = nr_rows * g.shape.nr_cols;
    CU(cudaMemcpyAsync(t->dev, src, n * 4, src_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, s));
    if (kind == EMB_TABLE_FIXED && src_f32) {
        uint32_t grid = (uint32_t)std::min<uint64_t>((n + 255) / 256, (uint64_t)g.sm_count * 32);
        emb::quantise_kernel<<<grid, 256, 0, s>>>((int32_t *)t->dev, (const float *)t->dev, n);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        CU(cudaGetLastError());
    }
    CU(cudaStreamSynchronize(s));
    return EMB_OK;
}

int emb_populate_table_i32(uint32_t table_id, uint64_t nr_rows, const int32_t *row_major)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    UseCtx on_owner(owner_of(table_id));
    return populate_whole(table_id, nr_rows, row_major, false, false, EMB_TABLE_FIXED, g.stream);
}

int emb_populate_table_f32(uint32_t table_id, uint64_t nr_rows, const float *row_major, int kind)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    UseCtx on_owner(owner_of(table_id));
    return populate_whole(table_id, nr_rows, row_major, false, true, kind, g.stream);
}

int emb_populate_table_device(uint32_t table_id, uint64_t nr_rows, const void *dev_row_major,
                              int src_is_f32, int kind, void *stream)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    UseCtx on_owner(owner_of(table_id));
    // in a device group the caller's stream belongs to another GPU than the table's for most tables: the copy then
    // runs on the owner's own stream and the source must be complete when the call is made
    return populate_whole(table_id, nr_rows, dev_row_major, true, src_is_f32 != 0, kind,
                          (stream && g_grp.n <= 1) ? (cudaStream_t)stream : g.stream);
}

int emb_table_set_row_base(uint32_t table_id, uint64_t row_base)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    UseCtx on_owner(owner_of(table_id));
    if (table_id >= g.tables.size() || !g.tables[table_id].kind)
        return fail(EMB_ESTATE, "table %u is not populated", table_id);
    if (row_base > 0xffffffffull)
        return fail(EMB_EINVAL, "row_base too large");
    if (g.tables[table_id].d_remap)
        return fail(EMB_EINVAL, "table %u is reordered (emb_table_remap): it cannot become a row shard", table_id);
    g.tables[table_id].row_base = row_base;
    g.tables[table_id].sharded = true;
    return EMB_OK;
}

int emb_table_read(uint32_t table_id, uint64_t row0, uint64_t n, void *dst_host)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    UseCtx on_owner(owner_of(table_id));
    if (table_id >= g.tables.size() || !g.tables[table_id].kind)
        return fail(EMB_ESTATE, "table %u is not populated", table_id);
    const Table &t = g.tables[table_id];
    if (row0 + n > t.rows || !dst_host)
        return fail(EMB_EINVAL, "rows [%llu, %llu) outside table of %llu rows", (unsigned long long)row0,
                    (unsigned long long)(row0 + n), (unsigned long long)t.rows);
    size_t rb = (size_t)g.shape.nr_cols * 4;
    CU(cudaMemcpy(dst_host, (const char *)t.dev + row0 * rb, n * rb, cudaMemcpyDeviceToHost));
    return EMB_OK;
}

int emb_table_info(uint32_t table_id, uint64_t *nr_rows, int *kind, const void **dev_ptr)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    UseCtx on_owner(owner_of(table_id));
    if (table_id >= g.tables.size() || !g.tables[table_id].kind)
        return fail(EMB_ESTATE, "table %u is not populated", table_id);
    const Table &t = g.tables[table_id];
    if (nr_rows)
        *nr_rows = t.rows;
    if (kind)
        *kind = t.kind;
    if (dev_ptr)
        *dev_ptr = t.dev;
    return EMB_OK;
}

int emb_free_table(uint32_t table_id)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    UseCtx on_owner(owner_of(table_id));
    if (table_id >= g.tables.size() || !g.tables[table_id].kind)
        return fail(EMB_ESTATE, "table %u is not populated", table_id);
    host_drain();
    CU(cudaDeviceSynchronize());
    cudaFree(g.tables[table_id].dev);
    cudaFree(g.tables[table_id].d_tmap);
    cudaFree(g.tables[table_id].d_remap);
    const uint32_t gen = g.tables[table_id].generation + 1;  // plans made for the freed slab refuse to launch
    g.tables[table_id] = Table();
    g.tables[table_id].generation = gen;
    return EMB_OK;
}

int emb_l2_persist(uint32_t table_id, uint64_t first_row, uint64_t nr_rows, float hit_ratio, void *stream)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    const int owner = nr_rows ? owner_of(table_id) : 0;
    if (owner != 0 && stream)
        return fail(EMB_EINVAL, "emb_l2_persist: table %u lives on GPU %d of the device group; a caller's stream can only "
                                "belong to the group's first GPU (pass NULL for the library's stream on the owner)",
                    table_id, owner);
    UseCtx on_owner(owner);
    cudaStream_t s = stream ? (cudaStream_t)stream : g.stream;
    cudaStreamAttrValue attr;
    memset(&attr, 0, sizeof attr);
    if (nr_rows == 0) {
        attr.accessPolicyWindow.num_bytes = 0;
        CU(cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &attr));
        CU(cudaCtxResetPersistingL2Cache());
        return EMB_OK;
    }
    if (table_id >= g.tables.size() || !g.tables[table_id].kind)
        return fail(EMB_ESTATE, "table %u is not populated", table_id);
    const Table &t = g.tables[table_id];
    if (first_row + nr_rows > t.rows)
        return fail(EMB_EINVAL, "window outside the table");
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, g.device));
    size_t rb = (size_t)g.shape.nr_cols * 4;
    size_t bytes = std::min((size_t)nr_rows * rb, (size_t)prop.accessPolicyMaxWindowSize);
    CU(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize,
                          std::min(bytes, (size_t)prop.persistingL2CacheMaxSize)));
    attr.accessPolicyWindow.base_ptr = (char *)t.dev + first_row * rb;
    attr.accessPolicyWindow.num_bytes = bytes;
    attr.accessPolicyWindow.hitRatio = hit_ratio;
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    CU(cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &attr));
    return EMB_OK;
}

/* ---------------------------------------------------------------- hot rows ------------- */

int emb_table_set_hot_rows(uint32_t table_id, uint64_t hot_rows, int mode)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    UseCtx on_owner(owner_of(table_id));
    if (table_id >= g.tables.size() || !g.tables[table_id].kind)
        return fail(EMB_ESTATE, "table %u is not populated", table_id);
    if (mode < 0 || mode > (EMB_HOT_KEEP | EMB_HOT_STREAM_COLD))
        return fail(EMB_EINVAL, "mode must be a combination of EMB_HOT_KEEP and EMB_HOT_STREAM_COLD");
    Table &t = g.tables[table_id];
    t.hot_rows = std::min<uint64_t>(hot_rows, t.rows);
    t.hot_mode = (uint32_t)mode;
    return EMB_OK;
}

int emb_hot_order(const uint64_t *counts, uint64_t nr_rows, uint32_t *new_of_old)
{
    if (!counts || !new_of_old || !nr_rows || nr_rows > 0xfffffffeull)
        return fail(EMB_EINVAL, "emb_hot_order: NULL arguments or nr_rows outside 1 .. 2^32-2");
    std::vector<uint32_t> order((size_t)nr_rows);
    for (uint64_t r = 0; r < nr_rows; r++)
        order[(size_t)r] = (uint32_t)r;
    // most looked-up rows first; rows with equal counts keep their order
    std::stable_sort(order.begin(), order.end(), [counts](uint32_t a, uint32_t b) { return counts[a] > counts[b]; });
    for (uint64_t k = 0; k < nr_rows; k++)
        new_of_old[order[(size_t)k]] = (uint32_t)k;
    return EMB_OK;
}

int emb_table_remap(uint32_t table_id, const uint32_t *new_of_old)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    UseCtx on_owner(owner_of(table_id));
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (table_id >= g.tables.size() || !g.tables[table_id].kind)
        return fail(EMB_ESTATE, "table %u is not populated", table_id);
    Table &t = g.tables[table_id];
    if (!new_of_old)
        return fail(EMB_EINVAL, "emb_table_remap: NULL permutation");
    if (t.sharded)
        return fail(EMB_EINVAL, "table %u is a row shard: reorder whole tables only", table_id);
    if (t.d_remap)
        return fail(EMB_ESTATE, "table %u is already reordered: populate it again first", table_id);
    // Only the stretch kernel translates indices (apply_hot): refuse now what every later lookup would refuse -- and
    // lookup() has no error return, it would end the process.
    if (!lanes_for_dim(g.shape.nr_cols))
        return fail(EMB_EINVAL, "emb_table_remap: NR_COLS %u is pooled by the any-dimension kernel, which cannot translate "
                                "indices (needs 4, 8, 16, 32, 64 or 128 columns)", g.shape.nr_cols);
    if (g_tune.variant == 1 || g_tune.variant == 4)
        return fail(EMB_EINVAL, "emb_table_remap: emb_set_tuning selected kernel variant %d, which cannot translate indices "
                                "(variant 6 or the default)", g_tune.variant);
    const uint64_t rows = t.rows;
    {  // a permutation of 0 .. rows-1?
        std::vector<bool> seen((size_t)rows, false);
        for (uint64_t r = 0; r < rows; r++) {
            const uint32_t v = new_of_old[r];
            if (v >= rows || seen[v])
                return fail(EMB_EINVAL, "emb_table_remap: new_of_old is not a permutation of 0..%llu (entry %llu = %u)",
                            (unsigned long long)(rows - 1), (unsigned long long)r, v);
            seen[v] = true;
        }
    }
    host_drain();
    CU(cudaDeviceSynchronize());
    const uint32_t D = g.shape.nr_cols;
    const size_t rb = (size_t)D * 4;
    uint32_t *d_map = nullptr;
    void *slab = nullptr;
    cudaError_t e = cudaMalloc(&d_map, (size_t)rows * 4);
    if (e == cudaSuccess)
        e = cudaMalloc(&slab, (size_t)(rows + 1) * rb);
    if (e != cudaSuccess) {
        cudaFree(d_map);
        return fail(EMB_ENOMEM, "emb_table_remap needs a second slab of %zu bytes while it reorders: %s",
                    (size_t)(rows + 1) * rb, cudaGetErrorString(e));
    }
    e = cudaMemcpyAsync(d_map, new_of_old, (size_t)rows * 4, cudaMemcpyHostToDevice, g.stream);
    if (e == cudaSuccess)
        e = cudaMemsetAsync((char *)slab + (size_t)rows * rb, 0, rb, g.stream);  // the all-zero row behind the table
    if (e == cudaSuccess) {
        const uint64_t n = rows * D;
        const uint32_t grid = (uint32_t)std::min<uint64_t>((n + 255) / 256, (uint64_t)g.sm_count * 32);
        emb::permute_rows_kernel<<<grid, 256, 0, g.stream>>>((uint32_t *)slab, (const uint32_t *)t.dev, d_map, rows, D);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(g.stream);  // new_of_old may be freed on return
    if (e != cudaSuccess) {
        cudaFree(d_map);
        cudaFree(slab);
        return fail(EMB_ECUDA, "emb_table_remap: %s", cudaGetErrorString(e));
    }
    cudaFree(t.dev);
    if (t.d_tmap) {
        cudaFree(t.d_tmap);
        t.d_tmap = nullptr;
    }
    t.dev = slab;
    t.d_remap = d_map;
    t.generation++;  // plans made for the old slab refuse to launch
    return EMB_OK;
}

int emb_remap_indices(uint32_t table_id, uint32_t *indices_dev, uint64_t n, void *stream)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    if (owner_of(table_id) != 0)
        return fail(EMB_EINVAL, "emb_remap_indices: table %u lives on GPU %d of the device group; calls that take device "
                                "pointers act on the group's first GPU only", table_id, owner_of(table_id));
    if (table_id >= g.tables.size() || !g.tables[table_id].kind)
        return fail(EMB_ESTATE, "table %u is not populated", table_id);
    const Table &t = g.tables[table_id];
    if (!t.d_remap || !n)
        return EMB_OK;  // rows are where the caller's ids say
    if (!indices_dev)
        return fail(EMB_EINVAL, "emb_remap_indices: NULL indices");
    cudaStream_t s = stream ? (cudaStream_t)stream : g.stream;
    const uint32_t grid = (uint32_t)std::min<uint64_t>((n + 255) / 256, (uint64_t)g.sm_count * 32);
    emb::remap_indices_kernel<<<grid, 256, 0, s>>>(indices_dev, t.d_remap, n, (uint32_t)t.rows);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());
    return EMB_OK;
}

/* ---------------------------------------------------------------- plans ---------------- */

int emb_plan_create(const emb_dev_query *queries, uint32_t n, emb_plan **out)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!queries || !n || !out)
        return fail(EMB_EINVAL, "emb_plan_create: NULL/empty arguments");
    emb_plan *p = new (std::nothrow) emb_plan();
    if (!p)
        return fail(EMB_ENOMEM, "out of host memory");
    p->h.resize(n);
    p->ctx = g_cur;
    const uint64_t D = g.shape.nr_cols;
    bool any_float = false, all_float = true, overlap = true;
    for (uint32_t i = 0; i < n; i++) {
        const bool f = queries[i].table_id < g.tables.size() && g.tables[queries[i].table_id].kind == EMB_TABLE_FLOAT;
        any_float |= f;
        all_float &= f;
        overlap &= (queries[i].flags & EMB_Q_OVERLAP) != 0;
    }
    p->all_float = all_float;
    p->overlap = overlap;
    {
        uint64_t li = 0, nb = 0;
        for (uint32_t i = 0; i < n; i++) {
            li += queries[i].indices_len;
            nb += queries[i].nr_batches;
        }
        p->variant = pick_variant(any_float, nb ? (double)li / (double)nb : 0.0, all_float, nb);
    }
    // Row shards of fixed-point tables: the compacting instantiation (K = 8).  EMB_SHARD_COMPACT=0 (measurement
    // knob) pools them with the plain kernel, where a foreign index gathers the all-zero row.
    static const bool compact_on = env_u32("EMB_SHARD_COMPACT", 1) != 0;
    if (compact_on && p->variant.kind == 6 && lanes_for_dim((uint32_t)D) && !any_float)
        for (uint32_t i = 0; i < n; i++)
            p->compact |= queries[i].table_id < g.tables.size() && g.tables[queries[i].table_id].sharded;
    if (p->compact && p->variant.unroll != 8) {
        p->variant.unroll = 8;
        if (p->variant.mult > 2)
            p->variant.mult = 2;
        p->variant.bags = std::min<uint32_t>(32 / lanes_for_dim((uint32_t)D) * p->variant.mult, 64);
    }
    {
        bool any_hot = false, any_remap = false;
        for (uint32_t i = 0; i < n; i++)
            if (queries[i].table_id < g.tables.size()) {
                const Table &tb = g.tables[queries[i].table_id];
                any_hot |= tb.hot_mode != 0 && tb.hot_rows != 0;
                any_remap |= tb.d_remap != nullptr && !(queries[i].flags & EMB_Q_REMAPPED);
            }
        if ((rc = apply_hot(p->variant, any_hot, any_remap, p->compact))) {
            delete p;
            return rc;
        }
    }
    uint64_t chunk = 0;
    for (uint32_t i = 0; i < n; i++) {
        uint32_t c = 0;
        if ((rc = fill_desc(queries[i], &p->h[i], (uint32_t)chunk, p->variant, &c))) {
            delete p;
            return rc;
        }
        chunk += c;
        p->gen.emplace_back(queries[i].table_id, g.tables[queries[i].table_id].generation);
        {  // a bag slice reads about its share of the indices (the host cannot see the device offsets)
            const uint64_t nbq = p->h[i].bag_end - p->h[i].bag_first;
            const uint64_t liq = queries[i].nr_batches ? (uint64_t)queries[i].indices_len * nbq / queries[i].nr_batches : 0;
            p->bytes += (liq + nbq) * (4 * D + 4);
        }
    }
    if (chunk > 0x7fffffffull) {
        delete p;
        return fail(EMB_EINVAL, "too many bags for one launch");
    }
    p->grid = (uint32_t)chunk;
    p->uni = uniform_chunks(p->h.data(), n);
    cudaError_t e = cudaMalloc(&p->d, n * sizeof(BagDesc));
    if (e == cudaSuccess)
        e = cudaMemcpy(p->d, p->h.data(), n * sizeof(BagDesc), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(p->d);
        delete p;
        return fail(EMB_ECUDA, "plan upload: %s", cudaGetErrorString(e));
    }
    *out = p;
    return EMB_OK;
}

int emb_plan_launch(emb_plan *plan, void *stream)
{
    if (!plan)
        return fail(EMB_EINVAL, "NULL plan");
    // the descriptors hold raw slab pointers: a table freed or re-allocated since the plan was made would be
    // read after free
    const Ctx &c = *plan->ctx;
    OnDevice on(c.device);
    for (const auto &tg : plan->gen)
        if (tg.first >= c.tables.size() || c.tables[tg.first].generation != tg.second || !c.tables[tg.first].kind)
            return fail(EMB_ESTATE, "plan is stale: table %u was freed or re-allocated after emb_plan_create", tg.first);
    return launch_bag_sum(c, plan->variant, plan->d, (uint32_t)plan->h.size(), plan->uni, plan->grid,
                          stream ? (cudaStream_t)stream : c.stream, plan->all_float, plan->overlap, plan->compact);
}

void emb_plan_destroy(emb_plan *plan)
{
    if (!plan)
        return;
    cudaFree(plan->d);
    delete plan;
}

uint64_t emb_plan_bytes(const emb_plan *plan) { return plan ? plan->bytes : 0; }
uint32_t emb_plan_grid(const emb_plan *plan) { return plan ? plan->grid : 0; }

/* One-shot, stream-ordered and asynchronous: the descriptors go through a small ring of
 * page-locked slots (each guarded by an event so a slot is not rewritten before its upload has
 * run), one cudaMemcpyAsync + one launch on the caller's stream, no allocation, no synchronise. */
int emb_lookup_device(const emb_dev_query *queries, uint32_t n, void *stream)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!queries || !n)
        return fail(EMB_EINVAL, "emb_lookup_device: NULL/empty arguments");
    cudaStream_t s = stream ? (cudaStream_t)stream : g.stream;
    if (g.ring_cap < n) {
        CU(cudaDeviceSynchronize());
        for (int i = 0; i < kRingSlots; i++) {
            if (g.ring_h[i])
                cudaFreeHost(g.ring_h[i]);
            if (g.ring_d[i])
                cudaFree(g.ring_d[i]);
            g.ring_h[i] = nullptr;
            g.ring_d[i] = nullptr;
        }
        g.ring_cap = 0;
        const uint32_t cap = std::max(n, 64u);
        for (int i = 0; i < kRingSlots; i++) {
            CU(cudaMallocHost(&g.ring_h[i], cap * sizeof(BagDesc)));
            CU(cudaMalloc(&g.ring_d[i], cap * sizeof(BagDesc)));
            if (!g.ring_ev[i])
                CU(cudaEventCreateWithFlags(&g.ring_ev[i], cudaEventDisableTiming));
        }
        g.ring_cap = cap;
    }
    const int slot = g.ring_next;
    g.ring_next = (g.ring_next + 1) % kRingSlots;
    CU(cudaEventSynchronize(g.ring_ev[slot]));  // the slot's previous upload AND the kernel reading it have executed
    BagDesc *h = g.ring_h[slot];
    bool any_float = false, all_float = true;
    for (uint32_t i = 0; i < n; i++) {
        const bool f = queries[i].table_id < g.tables.size() && g.tables[queries[i].table_id].kind == EMB_TABLE_FLOAT;
        any_float |= f;
        all_float &= f;
    }
    uint64_t tot_li = 0, tot_nb = 0;
    for (uint32_t i = 0; i < n; i++) {
        tot_li += queries[i].indices_len;
        tot_nb += queries[i].nr_batches;
    }
    Variant variant = pick_variant(any_float, tot_nb ? (double)tot_li / (double)tot_nb : 0.0, all_float, tot_nb);
    {
        bool any_hot = false, any_remap = false;
        for (uint32_t i = 0; i < n; i++)
            if (queries[i].table_id < g.tables.size()) {
                const Table &tb = g.tables[queries[i].table_id];
                any_hot |= tb.hot_mode != 0 && tb.hot_rows != 0;
                any_remap |= tb.d_remap != nullptr && !(queries[i].flags & EMB_Q_REMAPPED);
            }
        if ((rc = apply_hot(variant, any_hot, any_remap, false)))
            return rc;
    }
    uint64_t chunk = 0;
    for (uint32_t i = 0; i < n; i++) {
        uint32_t c = 0;
        if ((rc = fill_desc(queries[i], &h[i], (uint32_t)chunk, variant, &c)))
            return rc;
        chunk += c;
    }
    if (chunk > 0x7fffffffull)
        return fail(EMB_EINVAL, "too many bags for one launch");
    CU(cudaMemcpyAsync(g.ring_d[slot], h, n * sizeof(BagDesc), cudaMemcpyHostToDevice, s));
    rc = launch_bag_sum(g, variant, g.ring_d[slot], n, uniform_chunks(h, n), (uint32_t)chunk, s, all_float);
    // recorded AFTER the launch: the slot (pinned copy and device copy) is recycled only once the kernel that
    // reads its descriptors has run, whatever stream the next user of the slot launches on
    CU(cudaEventRecord(g.ring_ev[slot], s));
    return rc;
}

int emb_check_indices(void)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int any = 0;
    for (int k = 0; k < g_grp.n; k++) {  // every GPU of the device group
        UseCtx u(k);
        if (!g.device_ready)
            continue;
        uint32_t v = 0;
        CU(cudaDeviceSynchronize());
        CU(cudaMemcpy(&v, g.d_err, 4, cudaMemcpyDeviceToHost));
        if (v)
            CU(cudaMemset(g.d_err, 0, 4));
        any |= v ? 1 : 0;
    }
    return any;
}

int emb_reduce_partials(const void *const *parts_dev, uint32_t n_parts, uint32_t part_stride, void *out_dev,
                        uint32_t out_stride, uint64_t nr_bags, int is_raw_i32, void *stream)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!parts_dev || !n_parts || n_parts > EMB_MAX_PEERS || !out_dev)
        return fail(EMB_EINVAL, "emb_reduce_partials: bad arguments");
    if (!nr_bags)
        return EMB_OK;
    emb::PartList pl;
    memset(&pl, 0, sizeof pl);
    for (uint32_t s = 0; s < n_parts; s++)
        pl.p[s] = (const uint32_t *)parts_dev[s];
    uint32_t D = g.shape.nr_cols;
    uint64_t n = nr_bags * D;
    uint32_t grid = (uint32_t)std::min<uint64_t>((n + 255) / 256, (uint64_t)g.sm_count * 32);
    cudaStream_t s = stream ? (cudaStream_t)stream : g.stream;
    emb::reduce_partials_kernel<<<grid, 256, 0, s>>>(pl, n_parts, part_stride ? part_stride : D,
                                                     (uint32_t *)out_dev, out_stride ? out_stride : D, D,
                                                     nr_bags, is_raw_i32);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());
    return EMB_OK;
}

/* ---------------------------------------------------------------- exchange completion --- */

}  // extern "C"

struct emb_finish {
    emb::FlagList flags;
    uint32_t n_ranks = 0, my_rank = 0, n_jobs = 0, grid = 1;
    emb::ReduceJob *d_jobs = nullptr;
    Ctx *ctx = nullptr;  // where it was made (the group's first GPU); emb_finish_launch takes no lock, see emb_plan
};

extern "C" {

int emb_finish_create(void *const *flag_blocks, uint32_t n_ranks, uint32_t my_rank, const emb_reduce_job *jobs,
                      uint32_t n_jobs, emb_finish **out)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!flag_blocks || !n_ranks || n_ranks > EMB_MAX_PEERS || my_rank >= n_ranks || !out || (n_jobs && !jobs))
        return fail(EMB_EINVAL, "emb_finish_create: bad arguments");
    emb_finish *f = new (std::nothrow) emb_finish();
    if (!f)
        return fail(EMB_ENOMEM, "out of host memory");
    f->ctx = g_cur;
    memset(&f->flags, 0, sizeof f->flags);
    for (uint32_t r = 0; r < n_ranks; r++) {
        if (!flag_blocks[r]) {
            delete f;
            return fail(EMB_EINVAL, "flag_blocks[%u] is NULL", r);
        }
        f->flags.p[r] = (uint32_t *)flag_blocks[r];
    }
    f->n_ranks = n_ranks;
    f->my_rank = my_rank;
    f->n_jobs = n_jobs;
    const uint32_t D = g.shape.nr_cols;
    uint64_t work = 0;
    if (n_jobs) {
        std::vector<emb::ReduceJob> h(n_jobs);
        for (uint32_t j = 0; j < n_jobs; j++) {
            const emb_reduce_job &q = jobs[j];
            if (!q.n_parts || q.n_parts > EMB_MAX_PEERS || !q.out) {
                delete f;
                return fail(EMB_EINVAL, "reduce job %u: bad n_parts/out", j);
            }
            memset(&h[j], 0, sizeof h[j]);
            for (uint32_t s = 0; s < q.n_parts; s++) {
                if (!q.parts[s]) {
                    delete f;
                    return fail(EMB_EINVAL, "reduce job %u: parts[%u] is NULL", j, s);
                }
                h[j].parts[s] = (const uint32_t *)q.parts[s];
            }
            h[j].out = (uint32_t *)q.out;
            h[j].n_parts = q.n_parts;
            h[j].part_stride = q.part_stride ? q.part_stride : D;
            h[j].out_stride = q.out_stride ? q.out_stride : D;
            h[j].is_raw = q.is_raw_i32 ? 1u : 0u;
            h[j].nr_bags = q.nr_bags;
            work = std::max<uint64_t>(work, q.nr_bags * D);
        }
        cudaError_t e = cudaMalloc(&f->d_jobs, n_jobs * sizeof(emb::ReduceJob));
        if (e == cudaSuccess)
            e = cudaMemcpy(f->d_jobs, h.data(), n_jobs * sizeof(emb::ReduceJob), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            cudaFree(f->d_jobs);
            delete f;
            return fail(EMB_ECUDA, "emb_finish_create: %s", cudaGetErrorString(e));
        }
    }
    // every thread handles ~4 elements of the largest job; one block when there is nothing to reduce
    f->grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>((work + 1023) / 1024, (uint64_t)g.sm_count * 4));
    *out = f;
    return EMB_OK;
}

int emb_finish_launch(emb_finish *f, uint32_t epoch, void *stream)
{
    if (!f)
        return fail(EMB_EINVAL, "NULL finish object");
    const Ctx &c = *f->ctx;
    if (!c.device_ready)
        return fail(EMB_ESTATE, "emb_finish_launch after emb_shutdown()");
    OnDevice on(c.device);
    cudaStream_t s = stream ? (cudaStream_t)stream : c.stream;
    launch_dependent(emb::exchange_finish_kernel, dim3(f->grid, 1, 1), 256, s, f->flags, f->n_ranks, f->my_rank, epoch, f->d_jobs, f->n_jobs,
                     c.shape.nr_cols, c.d_err);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());
    return EMB_OK;
}

void emb_finish_destroy(emb_finish *f)
{
    if (!f)
        return;
    cudaFree(f->d_jobs);
    delete f;
}

/* ---------------------------------------------------------------- host-pointer lookups -- */

/* Shared body of lookup() and emb_lookup_ex().
 * emb_host.h:258-287 pushed indices/offsets/lengths to every column DPU (indices replicated
 * NR_COLS times), launched, then pulled every DPU's results (:305-321).  Here every byte
 * crosses PCIe once.  How it crosses depends on what the caller's arrays are:
 *
 *   inputs   staged : cudaMemcpyAsync host -> HBM arena on the copy-in stream (any host memory)
 *            mapped : the pooling kernel reads the caller's arrays over PCIe (page-locked only)
 *   outputs  staged : pooled rows to an HBM arena, cudaMemcpyAsync -> host on the copy-out stream
 *            mapped : the kernel's store epilogue writes final_results[t] over PCIe (page-locked)
 *
 * Measured on the pool's B200 (tools/host_path_probe.py, C2 bytes): the copy engines read host
 * memory at 55 GB/s but SM loads from mapped host memory reach only 35 GB/s (64-byte read
 * requests, tag-limited), while SM stores to mapped host memory (posted writes) reach 52 GB/s,
 * the same as a D2H copy, and need no extra CUDA calls.  So large calls stage the inputs and
 * map the outputs; the tables are cut into groups so that group g's kernel is writing results
 * to the host while group g+1's indices arrive -- both PCIe directions busy.  Small calls (the
 * reference's own shapes, a few hundred KB) map both: one descriptor upload, one launch, one
 * synchronise, because every CUDA call costs several microseconds of host time. */
constexpr uint64_t kPipelineMinBytes = 4ull << 20;
constexpr uint64_t kMapInputsMaxBytes = 1ull << 20;  // inputs are read in place below this
constexpr uint64_t kBounceMaxBytes = 1ull << 20;     // pageable calls up to this size go through the bounce arena ...
constexpr uint64_t kBouncePerTable = 128ull << 10;   // ... or up to this much per table (see bounce_limit)
constexpr uint64_t kBounceCeiling = 8ull << 20;
constexpr uint64_t kTinyBytes = 64ull << 10;         // below this even page-locked arrays are bounced

/* Pageable arrays cost the staged path 3 cudaMemcpyAsync per table, each ~5.5 us of host time because
 * the driver stages pageable memory synchronously (RM2 as shipped, upmem/r.sh:6-10: T=32, 1.5 MB per
 * call -> 628 us staged against 125 us on page-locked arrays; profiles/r1_small_calls.md).  One CPU
 * memcpy per array into the page-locked arena is cheaper until the bytes outweigh those calls. */
static uint64_t bounce_limit(uint32_t T)
{
    return std::min(kBounceCeiling, std::max(kBounceMaxBytes, (uint64_t)T * kBouncePerTable));
}
constexpr uint32_t kMaxGroups = 32;

/* Device-side alias of a host pointer when the GPU can address it directly (memory from
 * cudaHostAlloc / cudaHostRegister / emb_host_alloc / emb_host_register), else NULL. */
static void *mapped_alias(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return (a.type == cudaMemoryTypeHost) ? a.devicePointer : nullptr;
}

/* Start address of the CUDA allocation (cudaHostAlloc / cudaHostRegister range) that holds a host pointer, 0 for
 * memory the driver does not know (plain malloc).  One cudaMemcpyAsync may not span two page-locked allocations even
 * when they happen to be adjacent in the address space, so copies of neighbouring arrays are merged only inside one
 * allocation (or between two arrays of unregistered memory). */
static uintptr_t alloc_base(const void *p)
{
    typedef CUresult (*GetAttr)(void *, CUpointer_attribute, CUdeviceptr);
    static GetAttr fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        cudaDriverEntryPointQueryResult qr;
        void *f = nullptr;
        if (cudaGetDriverEntryPoint("cuPointerGetAttribute", &f, cudaEnableDefault, &qr) == cudaSuccess)
            fn = (GetAttr)f;
        else
            cudaGetLastError();
    }
    if (!fn)
        return (uintptr_t)p;  // cannot tell: every array is its own run
    CUdeviceptr base = 0;
    if (fn(&base, CU_POINTER_ATTRIBUTE_RANGE_START_ADDR, (CUdeviceptr)(uintptr_t)p) != CUDA_SUCCESS)
        return 0;
    return (uintptr_t)base;
}

/* Enqueue one host-pointer call on slot `hs`: everything up to, but not including, the wait for its last copy or
 * kernel.  hs.p describes the call for host_complete().  `ticket` != 0: an asynchronous call -- the end of its work
 * is marked with hs.ev_done, because by the time it is waited for the streams may hold the next call's work too. */
static int host_enqueue(HostSlot &hs, uint64_t ticket, bool overlapped, uint32_t T, const uint32_t *const *indices,
                        const uint32_t *const *offsets, const uint32_t *indices_len, const uint32_t *nr_batches,
                        float *const *final_results, const uint32_t *tids = nullptr)
{
    // the T arrays describe tables tids[0..T) of the current context (a device group hands every GPU its own
    // tables); NULL: tables 0..T-1
    double t_begin = now_s();
    hs.p.active = false;
    const uint32_t D = g.shape.nr_cols;
    auto tid = [tids](uint32_t t) { return tids ? tids[t] : t; };
    uint64_t tot_idx = 0, tot_bags = 0;
    bool any_float = false, all_float = true;
    for (uint32_t t = 0; t < T; t++) {
        if (tid(t) >= g.tables.size() || !g.tables[tid(t)].kind)
            return fail(EMB_ESTATE, "table %u is not populated", tid(t));
        if ((indices_len[t] && !indices[t]) || (nr_batches[t] && (!offsets[t] || !final_results[t])))
            return fail(EMB_EINVAL, "NULL array for table %u", tid(t));
        tot_idx += indices_len[t];
        tot_bags += nr_batches[t];
        any_float |= g.tables[tid(t)].kind == EMB_TABLE_FLOAT;
        all_float &= g.tables[tid(t)].kind == EMB_TABLE_FLOAT;
    }
    int rc;
    if (hs.desc_cap < T) {
        if (hs.d_descs)
            cudaFree(hs.d_descs);
        if (hs.h_descs)
            cudaFreeHost(hs.h_descs);
        hs.d_descs = nullptr;
        hs.h_descs = nullptr;
        hs.desc_cap = 0;
        hs.desc_shadow_ok = false;
        CU(cudaMalloc(&hs.d_descs, T * sizeof(BagDesc)));
        CU(cudaMallocHost(&hs.h_descs, T * sizeof(BagDesc)));
        hs.desc_cap = T;
    }
    const uint64_t h2d_payload = (tot_idx + tot_bags) * 4, d2h_payload = tot_bags * D * 4;

    // which arrays can the GPU address in place?
    std::vector<const uint32_t *> a_idx(T, nullptr), a_off(T, nullptr);
    std::vector<float *> a_out(T, nullptr);
    // A tiny call (the reference's mini-batch-1 inference) is bounced whatever the arrays
    // are: copying a few KB costs less host time than asking the driver about 3 T pointers.
    const bool tiny = g_tune.host_path == EMB_HOST_AUTO && h2d_payload + d2h_payload <= kTinyBytes;
    bool out_mapped = g_tune.host_path != EMB_HOST_COPY && !tiny, in_mapped = out_mapped;
    for (uint32_t t = 0; t < T && out_mapped; t++)
        if (nr_batches[t] && !(a_out[t] = (float *)mapped_alias(final_results[t])))
            out_mapped = false;
    in_mapped = out_mapped && (g_tune.host_path == EMB_HOST_MAPPED || h2d_payload <= kMapInputsMaxBytes);
    for (uint32_t t = 0; t < T && in_mapped; t++) {
        if (indices_len[t] && !(a_idx[t] = (const uint32_t *)mapped_alias(indices[t])))
            in_mapped = false;
        if (nr_batches[t] && !(a_off[t] = (const uint32_t *)mapped_alias(offsets[t])))
            in_mapped = false;
    }
    // Small call on arrays the GPU cannot address (plain malloc / NumPy memory -- what the
    // reference's callers pass): bounce through one page-locked arena.  The CPU copies a few
    // hundred KB faster than the 3 T cudaMemcpyAsync calls of the staged path cost to enqueue
    // (each CUDA call is several microseconds of host time), and the launch is the mapped one.
    bool bounce = false;
    if ((g_tune.host_path == EMB_HOST_AUTO && !(in_mapped && out_mapped) && h2d_payload + d2h_payload <= bounce_limit(T)) ||
        g_tune.host_path == EMB_HOST_BOUNCE) {
        const size_t need = (size_t)(h2d_payload + d2h_payload) + 64;
        if (hs.bounce_cap < need) {
            if (hs.h_bounce)
                cudaFreeHost(hs.h_bounce);
            hs.h_bounce = nullptr;
            hs.bounce_cap = 0;
            const size_t cap = std::max(need, (size_t)bounce_limit(T) + 64);
            CU(cudaHostAlloc((void **)&hs.h_bounce, cap, cudaHostAllocMapped | cudaHostAllocPortable));
            CU(cudaHostGetDevicePointer((void **)&hs.d_bounce, hs.h_bounce, 0));
            hs.bounce_cap = cap;
        }
        uint8_t *const dev_base = hs.d_bounce;
        size_t pos = 0;
        for (uint32_t t = 0; t < T; t++) {  // indices, then offsets (4-byte items), then 16-byte aligned outputs
            memcpy(hs.h_bounce + pos, indices[t], (size_t)indices_len[t] * 4);
            a_idx[t] = (const uint32_t *)(dev_base + pos);
            pos += (size_t)indices_len[t] * 4;
        }
        for (uint32_t t = 0; t < T; t++) {
            memcpy(hs.h_bounce + pos, offsets[t], (size_t)nr_batches[t] * 4);
            a_off[t] = (const uint32_t *)(dev_base + pos);
            pos += (size_t)nr_batches[t] * 4;
        }
        pos = (pos + 15) & ~(size_t)15;
        for (uint32_t t = 0; t < T; t++) {
            a_out[t] = (float *)(dev_base + pos);
            pos += (size_t)nr_batches[t] * D * 4;
        }
        bounce = in_mapped = out_mapped = true;
    }
    if (g_tune.host_path == EMB_HOST_MAPPED && !(in_mapped && out_mapped))
        return fail(EMB_EINVAL, "EMB_HOST_MAPPED: a host array is not page-locked (cudaHostAlloc / "
                                "cudaHostRegister / emb_host_alloc / emb_host_register)");
    if (g_tune.host_path == EMB_HOST_MAPPED_OUT && !out_mapped)
        return fail(EMB_EINVAL, "EMB_HOST_MAPPED_OUT: a final_results array is not page-locked");
    if (g_tune.host_path == EMB_HOST_MAPPED_OUT)
        in_mapped = false;
    // AUTO on a large call: both directions go through the copy engines, also when the result arrays are
    // page-locked.  Measured on C2 (tools/e2e_sweep.py, profiles/r2_host_path.md): 2.75 ms per call with copies
    // against 2.91 ms with the kernel's store epilogue writing over PCIe (arrays carved from one arena; about equal
    // with one allocation per table) -- a pooling kernel that stores to host memory is stretched to the PCIe
    // write rate and holds the SMs the next group's kernel wants.  EMB_AUTO_LARGE=0 (measurement knob) restores
    // the mapped results.
    static const bool auto_copy = env_u32("EMB_AUTO_LARGE", 1) != 0;
    if (g_tune.host_path == EMB_HOST_AUTO && auto_copy && out_mapped && !in_mapped && !bounce &&
        h2d_payload + d2h_payload >= kPipelineMinBytes)
        out_mapped = false;
    // MAPPED_OUT: the offsets (4 bytes per bag) are still read in place when they are page-locked:
    // a separate small copy per table costs more on the copy engine than it moves
    bool off_mapped = in_mapped;
    if (out_mapped && !in_mapped) {
        off_mapped = true;
        for (uint32_t t = 0; t < T && off_mapped; t++)
            if (nr_batches[t] && !(a_off[t] = (const uint32_t *)mapped_alias(offsets[t])))
                off_mapped = false;
    }

    if (!in_mapped && (rc = grow((void **)&hs.d_idx, &hs.idx_cap, tot_idx, 4)))
        return rc;
    if (!off_mapped && (rc = grow((void **)&hs.d_off, &hs.off_cap, tot_bags, 4)))
        return rc;
    if (!out_mapped && (rc = grow((void **)&hs.d_out, &hs.out_cap, tot_bags * D, 4)))
        return rc;

    // Groups of consecutive tables with about equal bytes (one group when nothing is staged).
    // A kernel that stores to host memory only retires when its writes have drained (measured,
    // tools/pcie_mix_probe.py: the same 109 MB take 2.08 ms in 1 launch, 2.13 in 4, 2.46 in 26) and
    // every copy boundary idles the copy engine for a few microseconds, so groups are large: ~32 MB.
    // Sweep on C2 (tools/e2e_sweep.py, arrays from one arena, 187 MB per call): copies both ways, equal groups,
    // 4 / 6 / 8 / 12 groups -> 2.80 / 2.75 / 2.76 / 2.95 ms; mapped results, ramped groups -> 2.91 / 2.92 / 2.98 /
    // 3.15 ms; floor of the two copies alone, running concurrently: 2.17 ms.
    const uint64_t staged = (in_mapped ? 0 : h2d_payload) + d2h_payload;
    uint32_t n_groups = 1;
    if (!(in_mapped && out_mapped) && staged >= kPipelineMinBytes) {
        const uint64_t unit = 32ull << 20;  // C2 (187 MB): 6 groups, the sweep's best for copies; 4-8 are within 1 %
        n_groups = std::min(std::min(T, kMaxGroups), std::max(2u, (uint32_t)((staged + unit / 2) / unit)));
    }
    const uint32_t forced_groups = env_u32("EMB_GROUPS", 0);  // measurement knob
    // `overlapped`: an asynchronous call issued while the previous one is still in flight.  The overlap that the
    // groups create inside one call is then there ACROSS calls (this call's copy-in runs under the previous call's
    // kernels and copy-out), and every group boundary only idles a copy engine: one group.  C2, two calls in flight
    // (tools/e2e_sweep.py --in-flight, profiles/r2_host_async.md): 1 / 2 / 3 / 6 groups -> 2.17 / 2.29 / 2.31 / 2.41 ms
    // per call; one call at a time wants the groups: 3.62 / 3.09 / 2.88 / 2.75 ms.
    if (overlapped && !forced_groups)
        n_groups = 1;
    if (forced_groups && n_groups > 1)
        n_groups = std::min(std::min(T, kMaxGroups), forced_groups);
    if ((rc = ensure_events(hs, 3 * n_groups + 2)))
        return rc;
    Variant variant = pick_variant(any_float, tot_bags ? (double)tot_idx / (double)tot_bags : 0.0, all_float, tot_bags);
    {  // the caller's indices are in the caller's row order: reordered tables are translated by the kernel
        bool any_hot = false, any_remap = false;
        for (uint32_t t = 0; t < T; t++) {
            any_hot |= g.tables[tid(t)].hot_mode != 0 && g.tables[tid(t)].hot_rows != 0;
            any_remap |= g.tables[tid(t)].d_remap != nullptr;
        }
        if ((rc = apply_hot(variant, any_hot, any_remap, false)))
            return rc;
    }
    std::vector<uint32_t> g_first(n_groups + 1, T), g_grid(n_groups, 0);
    {
        // Group sizes.  Nothing can be pooled before the first group's indices have crossed PCIe and nothing
        // overlaps the last group's results, so small groups at both ends shorten the two bubbles, while large
        // groups in the middle keep the copies long (every copy boundary idles the copy engine for a few
        // microseconds).  EMB_GROUP_SHAPE (measurement knob): 0 = equal groups (default when the results are copied:
        // 2.75 ms against 2.80 for the ramp on C2), 1 = sizes 1,2,4,...,4,2,1 (default with mapped results: the
        // small first kernel starts writing early), 2 = equal groups with a half-size last one.
        const uint64_t total = h2d_payload + d2h_payload;
        const uint32_t shape = env_u32("EMB_GROUP_SHAPE", out_mapped ? 1 : 0);
        std::vector<double> cum(n_groups);
        double sum = 0.0;
        for (uint32_t i = 0; i < n_groups; i++) {
            double w = 1.0;
            if (shape == 1)
                w = (double)std::min(std::min(1u << std::min(i, 16u), 1u << std::min(n_groups - 1 - i, 16u)), 4u);
            else if (shape == 2 && i + 1 == n_groups)
                w = 0.5;
            cum[i] = (sum += w);
        }
        uint64_t acc = 0;
        uint32_t gi = 0;
        g_first[0] = 0;
        for (uint32_t t = 0; t < T; t++) {
            acc += ((uint64_t)indices_len[t] + nr_batches[t]) * 4 + (uint64_t)nr_batches[t] * D * 4;
            if ((double)acc >= (double)total * cum[gi] / sum && gi + 1 < n_groups && t + 1 < T)
                g_first[++gi] = t + 1;
        }
        n_groups = gi + 1;
        g_first[n_groups] = T;
    }

    // descriptors: chunk_begin is relative to the group's launch
    uint64_t io = 0, oo = 0;
    std::vector<uint64_t> t_io(T), t_oo(T);
    for (uint32_t gi = 0; gi < n_groups; gi++) {
        uint64_t chunk = 0;
        for (uint32_t t = g_first[gi]; t < g_first[gi + 1]; t++) {
            emb_dev_query q;
            memset(&q, 0, sizeof q);
            q.table_id = tid(t);
            q.nr_batches = nr_batches[t];
            q.indices_len = indices_len[t];
            q.indices = in_mapped ? a_idx[t] : hs.d_idx + io;
            q.offsets = off_mapped ? a_off[t] : hs.d_off + oo;
            q.out = out_mapped ? (void *)a_out[t] : (void *)(hs.d_out + oo * D);
            q.out_stride = D;
            uint32_t c = 0;
            if ((rc = fill_desc(q, &hs.h_descs[t], (uint32_t)chunk, variant, &c)))
                return rc;
            chunk += c;
            t_io[t] = io;
            t_oo[t] = oo;
            io += indices_len[t];
            oo += nr_batches[t];
        }
        if (chunk > 0x7fffffffull)
            return fail(EMB_EINVAL, "too many bags for one launch");
        g_grid[gi] = (uint32_t)chunk;
    }

    // nothing to stage: descriptors and launch share one stream (fewer CUDA calls per small lookup)
    // Small calls are bound by host time per CUDA call and by the copy engine's start-up latency in
    // front of the launch, so nothing is enqueued that the call does not need: the descriptors are
    // uploaded only when they differ from what d_descs already holds (same arrays, same shapes, same
    // tables as the previous call -- the reference's caller reuses its buffers), and the begin/end
    // events exist only where a copy stream has something to time.
    cudaStream_t s_k = g.stream, s_in = in_mapped ? s_k : g.stream_in, s_out = g.stream_out;
    cudaEvent_t *ev = hs.ev_pool.data();
    cudaEvent_t ev_begin = ev[3 * n_groups], ev_end = ev[3 * n_groups + 1];
    if (s_in != s_k)
        CU(cudaEventRecord(ev_begin, s_in));
    static const bool desc_cache = env_u32("EMB_DESC_CACHE", 1) != 0;  // measurement knob
    const size_t desc_bytes = (size_t)T * sizeof(BagDesc);
    const bool desc_same = desc_cache && hs.desc_shadow_ok && hs.desc_shadow.size() == desc_bytes &&
                           memcmp(hs.desc_shadow.data(), hs.h_descs, desc_bytes) == 0;
    hs.desc_shadow_ok = false;  // set again once this call has completed
    if (!desc_same)
        CU(cudaMemcpyAsync(hs.d_descs, hs.h_descs, desc_bytes, cudaMemcpyHostToDevice, s_in));
    uint64_t launches = 0;
    for (uint32_t gi = 0; gi < n_groups; gi++) {
        const uint32_t t0 = g_first[gi], t1 = g_first[gi + 1];
        // Arrays of consecutive tables that are ADJACENT in host memory (a caller that carves its arrays out of one
        // arena) cross PCIe as one copy: the device arenas are contiguous in table order, so a run of adjacent host
        // arrays maps onto one contiguous device range.  Per-table copies cost the copy engine a start-up gap each
        // (52 + 26 copies: 66 GB/s both ways against 87 GB/s with one copy each way, profiles/r1_pcie.md).
        if (!in_mapped) {
            for (uint32_t t = t0; t < t1;) {  // indices
                uint32_t e = t + 1;
                size_t n = indices_len[t];
                const uintptr_t ab = alloc_base(indices[t]);
                while (e < t1 && indices[e] == indices[t] + n && alloc_base(indices[e]) == ab)
                    n += indices_len[e++];
                if (n)
                    CU(cudaMemcpyAsync(hs.d_idx + t_io[t], indices[t], n * 4, cudaMemcpyHostToDevice, s_in));
                t = e;
            }
            for (uint32_t t = t0; t < t1 && !off_mapped;) {  // offsets
                uint32_t e = t + 1;
                size_t n = nr_batches[t];
                const uintptr_t ab = alloc_base(offsets[t]);
                while (e < t1 && offsets[e] == offsets[t] + n && alloc_base(offsets[e]) == ab)
                    n += nr_batches[e++];
                if (n)
                    CU(cudaMemcpyAsync(hs.d_off + t_oo[t], offsets[t], n * 4, cudaMemcpyHostToDevice, s_in));
                t = e;
            }
        }
        if (s_in != s_k) {
            CU(cudaEventRecord(ev[3 * gi], s_in));
            CU(cudaStreamWaitEvent(s_k, ev[3 * gi], 0));
        }
        CU(cudaEventRecord(ev[3 * gi + 1], s_k));
        if ((rc = launch_bag_sum(g, variant, hs.d_descs + t0, t1 - t0, uniform_chunks(hs.h_descs + t0, t1 - t0), g_grid[gi], s_k,
                                 all_float)))
            return rc;
        launches += g_grid[gi] ? 1 : 0;
        CU(cudaEventRecord(ev[3 * gi + 2], s_k));
        if (!out_mapped) {
            CU(cudaStreamWaitEvent(s_out, ev[3 * gi + 2], 0));
            for (uint32_t t = t0; t < t1;) {  // adjacent result arrays: one copy
                uint32_t e = t + 1;
                size_t n = (size_t)nr_batches[t] * D;
                const uintptr_t ab = alloc_base(final_results[t]);
                while (e < t1 && final_results[e] == final_results[t] + n && alloc_base(final_results[e]) == ab)
                    n += (size_t)nr_batches[e++] * D;
                if (n)
                    CU(cudaMemcpyAsync(final_results[t], hs.d_out + t_oo[t] * D, n * 4, cudaMemcpyDeviceToHost, s_out));
                t = e;
            }
        }
    }
    cudaStream_t s_last = out_mapped ? s_k : s_out;
    if (!out_mapped)
        CU(cudaEventRecord(ev_end, s_last));
    if (ticket)
        CU(cudaEventRecord(hs.ev_done, s_last));  // created by ensure_events above
    HostPending &p = hs.p;
    p.ticket = ticket;
    p.t_begin = t_begin;
    p.T = T;
    p.n_groups = n_groups;
    p.timed_in = s_in != s_k;
    p.in_mapped = in_mapped;
    p.out_mapped = out_mapped;
    p.bounce = bounce;
    p.desc_same = desc_same;
    p.use_done_event = ticket != 0;
    p.s_last = s_last;
    p.h2d_payload = h2d_payload;
    p.d2h_payload = d2h_payload;
    p.launches = launches;
    p.desc_bytes = desc_bytes;
    p.b_dst.clear();
    p.b_src.clear();
    p.b_bytes.clear();
    if (bounce) {
        const uint8_t *dev_base = hs.d_bounce;
        for (uint32_t t = 0; t < T; t++) {
            p.b_dst.push_back(final_results[t]);
            p.b_src.push_back((size_t)((const uint8_t *)a_out[t] - dev_base));
            p.b_bytes.push_back((size_t)nr_batches[t] * D * 4);
        }
    }
    p.active = true;
    return EMB_OK;
}

/* Wait for the call enqueued on `hs`, then what only the host can do: stage timings, the bounce arena's copy-out. */
static int host_complete(HostSlot &hs)
{
    HostPending &p = hs.p;
    if (!p.active)
        return EMB_OK;
    p.active = false;
    if (p.use_done_event)
        CU(cudaEventSynchronize(hs.ev_done));
    else
        CU(cudaStreamSynchronize(p.s_last));
    if (!p.desc_same)
        hs.desc_shadow.assign((const uint8_t *)hs.h_descs, (const uint8_t *)hs.h_descs + p.desc_bytes);
    hs.desc_shadow_ok = true;
    cudaEvent_t *ev = hs.ev_pool.data();
    const uint32_t n_groups = p.n_groups;
    cudaEvent_t ev_begin = ev[3 * n_groups], ev_end = ev[3 * n_groups + 1];
    float ms = 0, k_ms = 0;
    g.last.h2d_ms = 0;
    if (p.timed_in) {
        CU(cudaEventElapsedTime(&ms, ev_begin, ev[3 * (n_groups - 1)]));
        g.last.h2d_ms = ms;
    }
    for (uint32_t gi = 0; gi < n_groups; gi++) {
        CU(cudaEventElapsedTime(&ms, ev[3 * gi + 1], ev[3 * gi + 2]));
        k_ms += ms;
    }
    g.last.kernel_ms = k_ms;  // with mapped arrays the PCIe traffic is inside the kernels
    g.last.d2h_ms = 0;
    if (!p.out_mapped) {
        CU(cudaEventElapsedTime(&ms, ev[2], ev_end));  // first group's rows ready -> last byte on the host
        g.last.d2h_ms = ms;
    }
    g.last.h2d_bytes = p.h2d_payload + p.T * sizeof(BagDesc);
    g.last.d2h_bytes = p.d2h_payload;
    g.last.launches = p.launches;
    if (p.bounce)
        for (uint32_t t = 0; t < p.T; t++)
            if (p.b_bytes[t])
                memcpy(p.b_dst[t], hs.h_bounce + p.b_src[t], p.b_bytes[t]);
    g.last.total_ms = (now_s() - p.t_begin) * 1e3;
    g.last.path = p.bounce ? EMB_HOST_BOUNCE
                  : (p.in_mapped && p.out_mapped) ? EMB_HOST_MAPPED : p.out_mapped ? EMB_HOST_MAPPED_OUT : EMB_HOST_COPY;
    return EMB_OK;
}

/* Calls of emb_lookup_ex_async still in flight complete before anything else touches the host path. */
static int host_drain()
{
    int rc = EMB_OK;
    for (HostSlot &hs : g.hslot) {
        const int r = host_complete(hs);
        if (r && !rc)
            rc = r;
    }
    return rc;
}

/* ... on every GPU of the device group. */
static int host_drain_group()
{
    int rc = EMB_OK;
    for (int k = 0; k < g_grp.n; k++) {
        UseCtx u(k);
        const int r = host_drain();
        if (r && !rc)
            rc = r;
    }
    return rc;
}

/* A call's tables cut by the GPU of the device group that holds them: whole tables, so the parts share nothing and
 * the reference's per-table result arrays need no exchange -- every GPU moves its tables' indices in and its tables'
 * pooled rows out over its own PCIe link, concurrently with the others. */
struct GroupParts {
    std::vector<const uint32_t *> idx[kMaxDevices], off[kMaxDevices];
    std::vector<float *> res[kMaxDevices];
    std::vector<uint32_t> ilen[kMaxDevices], nb[kMaxDevices], tids[kMaxDevices];

    GroupParts(uint32_t T, const uint32_t *const *indices, const uint32_t *const *offsets, const uint32_t *indices_len,
               const uint32_t *nr_batches, float *const *final_results)
    {
        for (uint32_t t = 0; t < T; t++) {
            const int k = owner_of(t);
            idx[k].push_back(indices[t]);
            off[k].push_back(offsets[t]);
            res[k].push_back(final_results[t]);
            ilen[k].push_back(indices_len[t]);
            nb[k].push_back(nr_batches[t]);
            tids[k].push_back(t);
        }
    }
    /* enqueue GPU k's part on its slot `slot` (the current context must be k) */
    int enqueue(int k, int slot, uint64_t ticket, bool overlapped)
    {
        return host_enqueue(g.hslot[slot], ticket, overlapped, (uint32_t)tids[k].size(), idx[k].data(), off[k].data(),
                            ilen[k].data(), nb[k].data(), res[k].data(), tids[k].data());
    }
};

/* emb_last_timings() of a call that ran on several GPUs: the stages overlap across GPUs, so times are the slowest
 * GPU's and bytes / launches are sums. */
static void merge_timings(emb_timings &a, const emb_timings &l)
{
    a.h2d_ms = std::max(a.h2d_ms, l.h2d_ms);
    a.kernel_ms = std::max(a.kernel_ms, l.kernel_ms);
    a.d2h_ms = std::max(a.d2h_ms, l.d2h_ms);
    a.total_ms = std::max(a.total_ms, l.total_ms);
    a.h2d_bytes += l.h2d_bytes;
    a.d2h_bytes += l.d2h_bytes;
    a.launches += l.launches;
    a.path = l.path;
}

static const emb_timings &last_call()
{
    return (g_grp.n > 1 && g_group_last_valid) ? g_group_last : g_ctxs[0].last;
}

/* The synchronous call: slot 0 (of every GPU of the group that holds one of the tables), enqueue + wait. */
static int lookup_host(uint32_t T, const uint32_t *const *indices, const uint32_t *const *offsets,
                       const uint32_t *indices_len, const uint32_t *nr_batches, float *const *final_results)
{
    int rc = host_drain_group();
    if (rc)
        return rc;
    if (g_grp.n <= 1) {
        if ((rc = host_enqueue(g.hslot[0], 0, false, T, indices, offsets, indices_len, nr_batches, final_results)))
            return rc;
        return host_complete(g.hslot[0]);
    }
    const double t0 = now_s();
    GroupParts gp(T, indices, offsets, indices_len, nr_batches, final_results);
    bool queued[kMaxDevices] = {};
    for (int k = 0; k < g_grp.n && !rc; k++) {  // everything is enqueued on every GPU before anything is waited for
        if (gp.tids[k].empty())
            continue;
        UseCtx u(k);
        if (!(rc = gp.enqueue(k, 0, 0, false)))
            queued[k] = true;
    }
    emb_timings agg = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int k = 0; k < g_grp.n; k++) {
        if (!queued[k])
            continue;
        UseCtx u(k);
        const int r = host_complete(g.hslot[0]);
        if (r && !rc)
            rc = r;
        merge_timings(agg, g.last);
    }
    agg.total_ms = (now_s() - t0) * 1e3;
    g_group_last = agg;
    g_group_last_valid = true;
    return rc;
}

int32_t *lookup(uint32_t **indices, uint32_t **offsets, float **final_results, void *dpu_set_ptr,
                int64_t latency_print)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        die("lookup", rc);
    if (dpu_set_ptr && dpu_set_ptr != (void *)&g_ctxs[0])
        fprintf(stderr, "emblib(B200): lookup: dpu_set_ptr %p is not the handle populate_mram returned (%p); ignored\n",
                dpu_set_ptr, (void *)&g_ctxs[0]);
    const uint32_t T = g.shape.nr_tables, B = g.shape.max_nr_batches;
    const uint32_t ilen = B * g.shape.max_indices_per_batch;  // INDICES_LEN, emb_host.h:30
    if (!B || !indices || !offsets || !final_results) {
        fail(EMB_EINVAL, "lookup: MAX_NR_BATCHES is 0 or NULL arguments");
        die("lookup", EMB_EINVAL);
    }
    std::vector<uint32_t> il(T, ilen), nb(T, B);  // struct query_len, emb_host.h:280-287
    rc = lookup_host(T, indices, offsets, il.data(), nb.data(), final_results);
    if (rc)
        die("lookup", rc);
    if (latency_print == 1) {  // the reference's six lines, emb_host.h:395-402
        const emb_timings &l = last_call();
        printf("C: Indices and offsets copying latency: %ldμs\n", (long)(l.h2d_ms * 1e3));
        printf("C: Query copying latency: %ldμs\n", 0L);
        printf("C: Dpu launch latency: %ldμs\n", (long)(l.kernel_ms * 1e3));
        printf("C: Results copy latency: %ldμs\n", (long)(l.d2h_ms * 1e3));
        printf("C: Callback prep latency: %ldμs\n", 0L);
        printf("C: DPU sync latency: %ldμs\n",
               (long)(std::max(0.0, l.total_ms - l.h2d_ms - l.kernel_ms - l.d2h_ms) * 1e3));
    }
    return nullptr;
}

int emb_lookup_ex(uint32_t nr_tables, const uint32_t *const *indices, const uint32_t *const *offsets,
                  const uint32_t *indices_len, const uint32_t *nr_batches, float *const *final_results)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!nr_tables || !indices || !offsets || !indices_len || !nr_batches || !final_results)
        return fail(EMB_EINVAL, "emb_lookup_ex: NULL/empty arguments");
    return lookup_host(nr_tables, indices, offsets, indices_len, nr_batches, final_results);
}

int emb_lookup_ex_async(uint32_t nr_tables, const uint32_t *const *indices, const uint32_t *const *offsets,
                        const uint32_t *indices_len, const uint32_t *nr_batches, float *const *final_results,
                        emb_ticket *ticket)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!nr_tables || !indices || !offsets || !indices_len || !nr_batches || !final_results || !ticket)
        return fail(EMB_EINVAL, "emb_lookup_ex_async: NULL/empty arguments");
    const uint64_t tk = g_next_ticket;
    const int slot = (int)(tk % kHostSlots), other = (int)((tk + 1) % kHostSlots);
    // the slot's previous call (ticket tk - 2) completes first if nobody has waited for it yet: its results are in
    // host memory from here on, and emb_lookup_wait on it returns at once
    bool overlapped = false;  // the previous call (ticket tk - 1) has not been waited for
    for (int k = 0; k < g_grp.n; k++) {
        UseCtx u(k);
        if ((rc = host_complete(g.hslot[slot])))
            return rc;
        overlapped |= g.hslot[other].p.active;
    }
    if (g_grp.n <= 1) {
        if ((rc = host_enqueue(g.hslot[slot], tk, overlapped, nr_tables, indices, offsets, indices_len, nr_batches,
                               final_results)))
            return rc;
    } else {
        GroupParts gp(nr_tables, indices, offsets, indices_len, nr_batches, final_results);
        for (int k = 0; k < g_grp.n; k++) {
            if (gp.tids[k].empty())
                continue;
            UseCtx u(k);
            if ((rc = gp.enqueue(k, slot, tk, overlapped)))
                return rc;  // parts already enqueued keep this ticket: the next call on the slot completes them
        }
    }
    g_next_ticket = tk + 1;
    *ticket = tk;
    return EMB_OK;
}

int emb_lookup_wait(emb_ticket ticket)
{
    std::unique_lock<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    if (!ticket || ticket >= g_next_ticket)
        return fail(EMB_EINVAL, "emb_lookup_wait: ticket %llu was never issued", (unsigned long long)ticket);
    const int slot = (int)(ticket % kHostSlots);
    emb_timings agg = {0, 0, 0, 0, 0, 0, 0, 0};
    bool any = false;
    int rc = EMB_OK;
    for (int k = 0; k < g_grp.n; k++) {  // the call's part on every GPU of the group
        HostSlot &hs = g_ctxs[k].hslot[slot];
        if (!hs.p.active || hs.p.ticket != ticket)
            continue;  // completed earlier (waited for already, or displaced by a later call on its slot)
        // Block outside the lock, so that another thread can enqueue the next call meanwhile.
        cudaEvent_t done = hs.ev_done;
        lk.unlock();
        cudaError_t e = cudaEventSynchronize(done);
        lk.lock();
        if (e != cudaSuccess)
            return fail(EMB_ECUDA, "emb_lookup_wait: %s", cudaGetErrorString(e));
        if (hs.p.active && hs.p.ticket == ticket) {
            UseCtx u(k);
            const int r = host_complete(hs);
            if (r && !rc)
                rc = r;
            merge_timings(agg, g.last);
            any = true;
        }
    }
    if (g_grp.n > 1 && any) {
        g_group_last = agg;
        g_group_last_valid = true;
    }
    return rc;
}

int emb_last_timings(emb_timings *out)
{
    if (!out)
        return fail(EMB_EINVAL, "NULL out");
    *out = last_call();
    return EMB_OK;
}

/* ---------------------------------------------------------------- memory helpers -------- */

int emb_device_alloc(void **dev_ptr, size_t nbytes)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!dev_ptr)
        return fail(EMB_EINVAL, "NULL dev_ptr");
    cudaError_t e = cudaMalloc(dev_ptr, nbytes ? nbytes : 1);
    if (e != cudaSuccess)
        return fail(EMB_ENOMEM, "cudaMalloc(%zu): %s", nbytes, cudaGetErrorString(e));
    return EMB_OK;
}

int emb_device_free(void *dev_ptr)
{
    CU(cudaFree(dev_ptr));
    return EMB_OK;
}

int emb_host_alloc(void **host_ptr, size_t nbytes)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!host_ptr)
        return fail(EMB_EINVAL, "NULL host_ptr");
    cudaError_t e = cudaMallocHost(host_ptr, nbytes ? nbytes : 1);
    if (e != cudaSuccess)
        return fail(EMB_ENOMEM, "cudaMallocHost(%zu): %s", nbytes, cudaGetErrorString(e));
    return EMB_OK;
}

int emb_host_free(void *host_ptr)
{
    CU(cudaFreeHost(host_ptr));
    return EMB_OK;
}

int emb_host_register(void *host_ptr, size_t nbytes)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DeviceScope caller_device;
    int rc = ensure_configured();
    if (rc)
        return rc;
    if (!host_ptr || !nbytes)
        return fail(EMB_EINVAL, "emb_host_register: NULL/empty range");
    CU(cudaHostRegister(host_ptr, nbytes, cudaHostRegisterMapped | cudaHostRegisterPortable));
    return EMB_OK;
}

int emb_host_unregister(void *host_ptr)
{
    CU(cudaHostUnregister(host_ptr));
    return EMB_OK;
}

/* These three take no lock: like every device-resident entry point they act on the group's first GPU, named
 * explicitly (the library's current context may be switched by a group call in another thread). */
int emb_memcpy_h2d(void *dst_dev, const void *src_host, size_t nbytes, void *stream)
{
    cudaStream_t s = stream ? (cudaStream_t)stream : g_ctxs[0].stream;
    CU(cudaMemcpyAsync(dst_dev, src_host, nbytes, cudaMemcpyHostToDevice, s));
    if (!stream)
        CU(cudaStreamSynchronize(s));
    return EMB_OK;
}

int emb_memcpy_d2h(void *dst_host, const void *src_dev, size_t nbytes, void *stream)
{
    cudaStream_t s = stream ? (cudaStream_t)stream : g_ctxs[0].stream;
    CU(cudaMemcpyAsync(dst_host, src_dev, nbytes, cudaMemcpyDeviceToHost, s));
    if (!stream)
        CU(cudaStreamSynchronize(s));
    return EMB_OK;
}

int emb_stream_sync(void *stream)
{
    if (stream)
        CU(cudaStreamSynchronize((cudaStream_t)stream));
    else if (g_ctxs[0].device_ready) {
        OnDevice on(g_ctxs[0].device);  // the library's GPU, whatever the caller's current device is
        CU(cudaDeviceSynchronize());
    }
    return EMB_OK;
}

int emb_ipc_export(void *dev_ptr, emb_ipc_handle *out)
{
    static_assert(sizeof(cudaIpcMemHandle_t) == EMB_IPC_HANDLE_BYTES, "IPC handle size");
    if (!dev_ptr || !out)
        return fail(EMB_EINVAL, "NULL argument");
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, dev_ptr));
    memcpy(out->bytes, &h, sizeof h);
    return EMB_OK;
}

int emb_ipc_open(const emb_ipc_handle *h, void **peer_dev_ptr)
{
    if (!h || !peer_dev_ptr)
        return fail(EMB_EINVAL, "NULL argument");
    cudaIpcMemHandle_t ch;
    memcpy(&ch, h->bytes, sizeof ch);
    CU(cudaIpcOpenMemHandle(peer_dev_ptr, ch, cudaIpcMemLazyEnablePeerAccess));
    return EMB_OK;
}

int emb_ipc_close(void *peer_dev_ptr)
{
    CU(cudaIpcCloseMemHandle(peer_dev_ptr));
    return EMB_OK;
}

}  // extern "C"

/* ---------------------------------------------------------------- pooling without any allocation --- */

/* emb_lookup_static (include/emb_b200.h): a lookup for callers that cannot let this library allocate.  Two kinds of
 * caller are meant.  One is a plugin inside a process whose own allocator has already taken (nearly) all of the GPU's
 * memory -- a serving engine or a framework's caching allocator that reserves HBM up front: every cudaMalloc of
 * ours (tables, the descriptor ring of emb_lookup_device, the error word) may fail there, or carve memory out from
 * under that allocator.  The other is a latency-bound loop: cudaMalloc and cudaFree synchronise the whole device, and
 * the other entry points may (re)allocate whenever a call is larger than the last one.
 *
 * The pooling kernels need two things beyond the caller's arrays: the list of per-table descriptors they read, and
 * the word an out-of-range index sets.  Everywhere else both live in cudaMalloc memory.  Here they live in this
 * file-scope array, which is part of the loaded module: every GPU that loads emblib.so holds these 11.5 KB from
 * the moment the module is loaded on it until the process exits, whether or not emb_lookup_static is ever called. */
namespace {

constexpr uint32_t kStaticMaxQueries = EMB_STATIC_MAX_QUERIES;
constexpr uint32_t kStaticPack = 16;  // descriptors per set-up launch: 16 * 176 B + 16 B of kernel parameters (< 4 KB)

struct StaticScratch {
    uint32_t err;  // bit 0: an index fell outside its table (cleared by every call before its pooling launch)
    uint32_t pad_[63];
    BagDesc descs[kStaticMaxQueries];
};
static_assert(sizeof(StaticScratch) == 256 + kStaticMaxQueries * sizeof(BagDesc), "static scratch layout");

static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

struct StaticPack {
    BagDesc d[kStaticPack];
};

/* Device code may name the array; host code may not (there the name is only a handle for cudaGetSymbolAddress).
 * Writes descriptors [first, first + count) from the kernel's own parameters -- no host buffer has to outlive the
 * call, no copy is staged -- and, in the first launch of a call, clears the error word. */
__global__ void static_setup_kernel(StaticPack pack, uint32_t first, uint32_t count, uint32_t clear_err)
{
    StaticScratch *sc = reinterpret_cast<StaticScratch *>(g_static_scratch);
    const uint4 *src = reinterpret_cast<const uint4 *>(pack.d);
    uint4 *dst = reinterpret_cast<uint4 *>(sc->descs + first);
    const uint32_t n16 = count * (uint32_t)(sizeof(BagDesc) / 16);
    for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x)
        dst[i] = src[i];
    if (clear_err && threadIdx.x == 0)
        sc->err = 0;
}

struct StaticDevice {
    bool ready = false;
    StaticScratch *scratch = nullptr;  // DEVICE address of g_static_scratch on this GPU
};
constexpr int kStaticMaxDevices = 64;
StaticDevice g_static_dev[kStaticMaxDevices];
std::mutex g_static_mu;

template <int LANES>
void launch_static(int unroll, uint32_t grid, cudaStream_t s, const BagDesc *d, uint32_t n, uint32_t uni, uint32_t *err)
{
    if (unroll == 8)
        emb::bag_sum_kernel<LANES, 8><<<grid, emb::kThreads, 0, s>>>(d, n, uni, err);
    else
        emb::bag_sum_kernel<LANES, 4><<<grid, emb::kThreads, 0, s>>>(d, n, uni, err);
}

}  // namespace

extern "C" {

int emb_lookup_static(const emb_static_query *queries, uint32_t n, uint32_t nr_cols, void *stream, int *out_of_range)
{
    if (out_of_range)
        *out_of_range = 0;
    if (!queries || !n)
        return fail(EMB_EINVAL, "emb_lookup_static: NULL/empty arguments");
    if (n > kStaticMaxQueries)
        return fail(EMB_EINVAL, "emb_lookup_static: %u queries in one call, the static descriptor array holds at most %u "
                                "(EMB_STATIC_MAX_QUERIES): split the call", n, kStaticMaxQueries);
    if (nr_cols == 0 || nr_cols > 4096)
        return fail(EMB_EINVAL, "emb_lookup_static: nr_cols %u unsupported (1 .. 4096)", nr_cols);
    const uint32_t D = nr_cols;
    const uint32_t lanes = lanes_for_dim(D);           // 0: the any-dimension kernel, one warp per bag
    const uint32_t per = emb::kThreads / (lanes ? lanes : 32);  // bags per CTA
    const int unroll = lanes >= 16 ? 8 : 4;

    // descriptors, built in ordinary host memory (they travel as kernel parameters below)
    BagDesc h[kStaticMaxQueries];
    uint64_t chunk = 0;
    for (uint32_t i = 0; i < n; i++) {
        const emb_static_query &q = queries[i];
        BagDesc *d = &h[i];
        memset(d, 0, sizeof *d);
        if (q.kind != EMB_TABLE_FIXED && q.kind != EMB_TABLE_FLOAT)
            return fail(EMB_EINVAL, "emb_lookup_static: query %u: kind %u is neither EMB_TABLE_FIXED nor EMB_TABLE_FLOAT", i, q.kind);
        if (!q.table || q.nr_rows == 0 || q.nr_rows > 0xfffffffeull)
            return fail(EMB_EINVAL, "emb_lookup_static: query %u: NULL table or nr_rows %llu outside 1 .. 2^32-2", i,
                        (unsigned long long)q.nr_rows);
        if (lanes && (uintptr_t)q.table % 16 != 0)
            return fail(EMB_EINVAL, "emb_lookup_static: query %u: table is not 16-byte aligned", i);
        if (q.flags & ~EMB_Q_RAW_SUMS)
            return fail(EMB_EINVAL, "emb_lookup_static: query %u: only EMB_Q_RAW_SUMS is understood here", i);
        const uint32_t stride = q.out_stride ? q.out_stride : D;
        if (stride < D)
            return fail(EMB_EINVAL, "emb_lookup_static: query %u: out_stride %u < nr_cols %u", i, stride, D);
        if (q.nr_batches && (!q.offsets || !q.out || (q.indices_len && !q.indices)))
            return fail(EMB_EINVAL, "emb_lookup_static: query %u has NULL indices/offsets/out", i);
        d->table = q.table;
        d->indices = q.indices;
        d->offsets = q.offsets;
        d->out = q.out;
        d->rows = (uint32_t)q.nr_rows;
        d->nr_batches = q.nr_batches;
        d->indices_len = q.indices_len;
        d->out_stride = stride;
        d->bag_first = 0;
        d->bag_end = q.nr_batches;
        uint32_t flags = 0;
        if (q.kind == EMB_TABLE_FLOAT)
            flags |= emb::kDescFloat;
        else if (q.flags & EMB_Q_RAW_SUMS)
            flags |= emb::kDescRaw;
        if (stride % 4 == 0 && (uintptr_t)q.out % 16 == 0)
            flags |= emb::kDescVecStore;
        d->flags = flags;
        d->chunk_begin = (uint32_t)chunk;
        d->n_chunks = (q.nr_batches + per - 1) / per;
        chunk += d->n_chunks;
        if (chunk > 0x7fffffffull)
            return fail(EMB_EINVAL, "emb_lookup_static: too many bags for one launch");
    }
    if (chunk == 0)
        return EMB_OK;  // no bags anywhere: nothing to pool, nothing to flag

    // One call at a time builds its descriptors in the shared array.  The lock orders the ENQUEUES of two threads; it
    // does not stop two calls on different streams from running on the GPU at the same time (see the header).
    std::lock_guard<std::mutex> lk(g_static_mu);
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kStaticMaxDevices)
        return fail(EMB_EINVAL, "emb_lookup_static: current device %d outside 0 .. %d", dev, kStaticMaxDevices - 1);
    StaticDevice &sd = g_static_dev[dev];
    if (!sd.ready) {
        // Set-up, once per GPU, with that GPU current: ask the runtime where THIS GPU keeps its copy of the array.
        // `g_static_scratch` itself is not that address in host code -- handing it to a kernel would make the kernel
        // fault or, where the GPU can reach host memory, quietly use the wrong bytes.
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p)
            return fail(EMB_ECUDA, "emb_lookup_static: cudaGetSymbolAddress(static scratch) on device %d -> %s", dev,
                        cudaGetErrorString(e));
        sd.scratch = static_cast<StaticScratch *>(p);
        sd.ready = true;
    }
    // NULL = the legacy default stream here: this entry point creates nothing, so it has no stream of its own
    cudaStream_t s = (cudaStream_t)stream;

    for (uint32_t first = 0; first < n; first += kStaticPack) {
        StaticPack pack;
        const uint32_t count = std::min(kStaticPack, n - first);
        memset(&pack, 0, sizeof pack);
        memcpy(pack.d, h + first, count * sizeof(BagDesc));
        static_setup_kernel<<<1, 128, 0, s>>>(pack, first, count, first == 0 ? 1u : 0u);
        g_launches.fetch_add(1, std::memory_order_relaxed);
    }
    CU(cudaGetLastError());

    const BagDesc *d_descs = sd.scratch->descs;  // address arithmetic only: nothing is dereferenced on the host
    uint32_t *d_err = &sd.scratch->err;
    const uint32_t uni = uniform_chunks(h, n);
    const uint32_t grid = (uint32_t)chunk;
    switch (lanes) {
    case 0: emb::bag_sum_generic_kernel<<<grid, emb::kThreads, 0, s>>>(d_descs, n, uni, D, d_err); break;
    case 1: launch_static<1>(unroll, grid, s, d_descs, n, uni, d_err); break;
    case 2: launch_static<2>(unroll, grid, s, d_descs, n, uni, d_err); break;
    case 4: launch_static<4>(unroll, grid, s, d_descs, n, uni, d_err); break;
    case 8: launch_static<8>(unroll, grid, s, d_descs, n, uni, d_err); break;
    case 16: launch_static<16>(unroll, grid, s, d_descs, n, uni, d_err); break;
    default: launch_static<32>(unroll, grid, s, d_descs, n, uni, d_err); break;
    }
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());

    if (out_of_range) {
        uint32_t v = 0;
        CU(cudaMemcpyAsync(&v, d_err, sizeof v, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        *out_of_range = (v & 1u) ? 1 : 0;
    }
    return EMB_OK;
}

}  // extern "C"
