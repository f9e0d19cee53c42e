/*
 * emb_kernels.cuh -- sm_100a kernels of the embedding-lookup hot path.
 *
 * What the UPMEM build spread over NR_TABLES*NR_COLS DPUs x 14 tasklets
 * (upmem/src/dpu/emb_dpu_lookup.c:106-128: one int32 column per DPU, one dependent 8-byte
 * MRAM read per index) is ONE launch here: tables are row-major [rows][D] in HBM, a
 * sub-warp of D/4 lanes owns a bag, every lane moves 16 bytes of every row of the bag
 * (LDG.128, the row is one or more fully used 32-byte sectors), and the reference's
 * post_process() (upmem/include/emb_host.h:207-212) is the store epilogue.
 *
 * HBM-bound integer work: no tensor cores, no shared-memory tiling of table data (no reuse);
 * the CTA stages only its bags' offsets and indices in shared memory so that the row
 * gathers of a sub-warp are independent loads that can all be in flight together.
 */
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/emb_b200.h"

namespace emb {

constexpr int kThreads = 256;    // CTA size: 8 warps
#ifndef EMB_IDX_STAGE
#define EMB_IDX_STAGE 3072
#endif
constexpr int kIdxStage = EMB_IDX_STAGE;  // indices staged per CTA and pass

enum : uint32_t {
    kDescFloat = 1u,    // table holds fp32
    kDescRaw = 2u,      // write raw int32 sums (no dequantise)
    kDescSharded = 4u,  // row-range shard: out-of-range indices are silently skipped
    kDescVecStore = 8u, // output rows are 16-byte aligned
};
enum : uint32_t {  // BagDesc::hot_mode
    kHotKeep = 1u,        // rows below hot_rows: L2 evict_last
    kHotStreamCold = 2u,  // rows from hot_rows on: L2 evict_first
};

/* One per query; lives in device memory for the lifetime of a plan. */
struct __align__(16) BagDesc {
    const void *table;        // [rows][D] int32 or fp32
    const uint32_t *indices;  // [indices_len]
    const uint32_t *offsets;  // [nr_batches]
    void *out;                // single-slab output (bags_per_rank == 0)
    void *out_peer[EMB_MAX_PEERS];
    uint32_t rows;       // rows held in this slot
    uint32_t row_base;   // global id of row 0
    uint32_t nr_batches;
    uint32_t indices_len;
    uint32_t out_stride;  // elements
    uint32_t bags_per_rank;
    uint32_t flags;
    uint32_t chunk_begin;  // first CTA of this query
    uint32_t bag_first;    // the query covers bags [bag_first, bag_end) of the nr_batches in the arrays
    uint32_t bag_end;
    uint32_t n_chunks;     // CTAs of this query
    uint32_t chunk_rot;    // CTA c pools chunk (c + chunk_rot) mod n_chunks: ranks start at different bags
    const void *tmap;      // CUtensorMap of the table in device memory (TMA kernel only), 64-byte aligned
    // Hot-row tables (emb_table_set_hot_rows / emb_table_remap; stretch kernel, HOT instantiation).  remap != NULL:
    // the slab was reordered, a caller's row id r lives at slab row remap[r] (translated while the indices are staged).
    // Slab rows below hot_rows are loaded with the L2 policy hot_mode bit 0 selects (evict_last), the others with the
    // one bit 1 selects (evict_first); a clear bit means evict_normal.
    const uint32_t *remap;
    uint32_t hot_rows;
    uint32_t hot_mode;
    uint64_t pad_;
};
static_assert(sizeof(BagDesc) == 176, "BagDesc layout");

/* First bag of the chunk CTA `cta` (relative to the query) pools; `per` = bags per CTA. */
__device__ __forceinline__ uint32_t chunk_first_bag(const BagDesc *d, uint32_t cta, uint32_t per)
{
    uint32_t c = cta + d->chunk_rot;
    if (c >= d->n_chunks)
        c -= d->n_chunks;
    return d->bag_first + c * per;
}

/* Row loads are plain ld.global.nc (L1 allocates, L2 default priority).  Measured in round 1 against
 * L1::no_allocate, L2 evict_last / evict_normal policy operands and the .L2::64B / .L2::128B prefetch sizes
 * (profiles/r1_sweeps.md): ncu counts every sector of an L1::no_allocate load as evict_first at L2, so the rows
 * of the table being walked were the first thing L2 dropped; plain ld.global.nc is faster on every workload
 * (C2 0.2443 -> 0.2422 ms, C3 0.0496 -> 0.0456, C4 0.658 -> 0.6088, C5 0.5023 -> 0.4819).  The index / offset
 * streams and the output rows stay L1::no_allocate: they are used once. */
/* One 16-byte piece of a table row through the read-only path.  NARROW (rows of 64 bytes or less) adds the L2
 * prefetch-size hint of 64 bytes: by default an L2 miss fills the whole 128-byte line, so a gather of 64-byte rows
 * reads twice its bytes from DRAM (C2, ncu dram__bytes_read per launch: 1.329 GB without the hint, 0.909 GB with it;
 * profiles/r2_row_loads.md).  The kernel's time does not change (0.2215 ms either way: it is not bound by DRAM
 * bytes), the traffic does.  Table slabs come from cudaMalloc (256-byte aligned), so a 64-byte row never straddles a
 * 64-byte boundary. */
template <bool NARROW>
__device__ __forceinline__ uint4 ldg_row16(const uint4 *p)
{
    uint4 r;
    if (NARROW)
        asm volatile("ld.global.nc.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    else
        asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

/* The same load with an L2 cache policy chosen per row (hot-row tables): `hot` picks pol_hot, else pol_cold.  Two
 * predicated loads into the same registers: no 64-bit select in front of the load. */
template <bool NARROW>
__device__ __forceinline__ uint4 ldg_row16_policy(const uint4 *p, bool hot, uint64_t pol_hot, uint64_t pol_cold)
{
    uint4 r;
    if (NARROW)
        asm volatile("{\n\t"
                     ".reg .pred h;\n\t"
                     "setp.ne.u32 h, %5, 0;\n\t"
                     "@h ld.global.nc.L2::cache_hint.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4], %6;\n\t"
                     "@!h ld.global.nc.L2::cache_hint.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4], %7;\n\t"
                     "}"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                     : "l"(p), "r"((uint32_t)hot), "l"(pol_hot), "l"(pol_cold));
    else
        asm volatile("{\n\t"
                     ".reg .pred h;\n\t"
                     "setp.ne.u32 h, %5, 0;\n\t"
                     "@h ld.global.nc.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %6;\n\t"
                     "@!h ld.global.nc.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %7;\n\t"
                     "}"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                     : "l"(p), "r"((uint32_t)hot), "l"(pol_hot), "l"(pol_cold));
    return r;
}
/* L2 cache policies for every byte of an access (createpolicy ... 1.0): 0 = evict_normal, 1 = evict_last,
 * 2 = evict_first. */
__device__ __forceinline__ uint64_t l2_policy(int kind)
{
    uint64_t pol;
    if (kind == 1)
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    else if (kind == 2)
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    else
        asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

__device__ __forceinline__ uint32_t ldg_stream_u32(const uint32_t *p)
{
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
/* The partial sums a peer stored into this rank's slab are read once, AFTER an acquire on the peer's flag: the
 * coherent path, not ld.global.nc. */
__device__ __forceinline__ uint4 ldg_stream16(const uint4 *p)
{
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

/* Result rows: written once, never read by this kernel.  Cache-streaming (.cs), write-through (.wt), plain stores
 * and L2 evict_first / evict_last policies all measure the same to four digits on both result layouts
 * (profiles/r2_scaling.md). */
__device__ __forceinline__ void stg_stream16(void *p, uint4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}

/* emb_host.h:210  (float)s / pow(10,9): int->float RNE, divide in double, round to float.
 * 1e9 is exact in fp32 and double has >= 2*24+2 bits, so this equals one IEEE fp32 divide x / 1e9f.
 * That divide is computed here without the division routine: with r = RN(1/1e9), q0 = RN(x*r),
 * e = x - q0*1e9 (exact in one FMA), q = RN(q0 + e*r) is the correctly rounded quotient (Markstein's
 * correction step).  Checked exhaustively on the CPU against `/` for every one of the 2^32 int32 inputs
 * (tests/check_dequant.c, run by tests/test_dequant_exhaustive.py): 0 mismatches.  4 instructions instead of ~12
 * and no slow-path branch. */
__device__ __forceinline__ float dequant(uint32_t s)
{
    const float x = __int2float_rn((int)s);
    const float r = 1.0f / 1.0e9f;  // folded at compile time, round to nearest: 0x3089705f
    const float q0 = __fmul_rn(x, r);
    const float e = __fmaf_rn(-q0, 1.0e9f, x);
    return __fmaf_rn(e, r, q0);
}

/* dlrm_dpu_pytorch.py:318 + :329: float32 product, rint (half-to-even), int64, low 32 bits. */
__device__ __forceinline__ int32_t quantise(float w)
{
    float r = rintf(__fmul_rn(w, 1.0e9f));
    // numpy's float->int64 cast of an out-of-range / NaN value is the x86 "integer
    // indefinite" 0x8000000000000000, whose low word is 0
    if (!(fabsf(r) < 9.2233720e18f))
        return 0;
    return (int32_t)(uint32_t)(unsigned long long)__float2ll_rn(r);
}

/* Descriptor of CTA `chunk`.  uni != 0: every descriptor owns `uni` consecutive CTAs (the host checked,
 * uniform_chunks()), so this is a division instead of a chain of log2(n) dependent loads in front of
 * everything else the CTA does. */
__device__ __forceinline__ uint32_t find_desc(const BagDesc *__restrict__ descs, uint32_t n, uint32_t uni,
                                              uint32_t chunk)
{
    if (uni)
        return chunk / uni;
    uint32_t lo = 0, hi = n;  // last desc with chunk_begin <= chunk
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (__ldg(&descs[mid].chunk_begin) <= chunk)
            lo = mid;
        else
            hi = mid;
    }
    return lo;
}

/*
 * Sum-pool kernel, D = 4*LANES columns (LANES in 1..32, power of two).
 * grid.x = sum over queries of ceil(nr_batches / (kThreads/LANES)); one CTA handles
 * kThreads/LANES consecutive bags of one table.
 */
template <int LANES, int UNROLL>
__global__ void __launch_bounds__(kThreads)
bag_sum_kernel(const BagDesc *__restrict__ descs, uint32_t n_desc, uint32_t uni, uint32_t *__restrict__ err_flag)
{
    constexpr int BAGS = kThreads / LANES;
    __shared__ uint32_t s_off[BAGS + 1];
    __shared__ uint32_t s_idx[kIdxStage];

    const uint32_t tid = threadIdx.x;
    const uint32_t q = find_desc(descs, n_desc, uni, blockIdx.x);
    const BagDesc *__restrict__ d = descs + q;

    const uint32_t B = d->nr_batches;
    const uint32_t ilen = d->indices_len;
    const uint32_t b0 = chunk_first_bag(d, blockIdx.x - d->chunk_begin, BAGS);
    const uint32_t nb = min((uint32_t)BAGS, d->bag_end - b0);
    const uint32_t *__restrict__ g_idx = d->indices;
    const uint32_t *__restrict__ g_off = d->offsets;

    // bag bounds, emb_dpu_lookup.c:60-63,109-110: bag 0 starts at 0 whatever offsets[0] is,
    // bag i ends where bag i+1 starts, the last bag ends at indices_len
    for (uint32_t i = tid; i <= nb; i += kThreads) {
        uint32_t b = b0 + i;
        uint32_t v = (b == 0) ? 0u : (b >= B ? ilen : ldg_stream_u32(g_off + b));
        s_off[i] = min(v, ilen);
    }
    __syncthreads();

    const uint32_t wlo = s_off[0], whi = s_off[nb];
    const bool staged = whi > wlo && (whi - wlo) <= (uint32_t)kIdxStage;
    if (staged) {
        const uint32_t n = whi - wlo;
        for (uint32_t i = tid; i < n; i += kThreads)
            s_idx[i] = ldg_stream_u32(g_idx + wlo + i);
    }
    __syncthreads();

    const uint32_t bag = tid / LANES;
    const uint32_t lane = tid % LANES;
    if (bag >= nb)
        return;

    const uint32_t start = s_off[bag];
    const uint32_t end = s_off[bag + 1];
    const uint32_t flags = d->flags;
    const uint32_t rows = d->rows;
    const uint32_t row_base = d->row_base;
    const uint4 *__restrict__ table = reinterpret_cast<const uint4 *>(d->table) + lane;
    const bool in_window = staged && start >= wlo && end <= whi;  // false only for non-monotone offsets

    uint4 acc = make_uint4(0u, 0u, 0u, 0u);
    float4 facc = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool is_float = flags & kDescFloat;
    bool bad = false;

    for (uint32_t p = start; p < end; p += UNROLL) {
        uint32_t r[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            uint32_t id = 0xffffffffu;
            if (p + u < end) {
                id = in_window ? s_idx[p + u - wlo] : __ldg(g_idx + p + u);
                id -= row_base;  // unsigned: rows below the shard wrap to huge values
                bad |= id >= rows;
            } else {
                id = 0xffffffffu;
            }
            r[u] = id;
        }
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            v[u] = make_uint4(0u, 0u, 0u, 0u);
            if (r[u] < rows)
                v[u] = ldg_row16<(LANES <= 4)>(table + (size_t)r[u] * LANES);
        }
        if (!is_float) {
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {  // int32 wrap-around adds (emb_dpu_lookup.c:114)
                acc.x += v[u].x;
                acc.y += v[u].y;
                acc.z += v[u].z;
                acc.w += v[u].w;
            }
        } else {
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {  // fp32, strictly in index order (embedding_lookup_idx.cc:33-77)
                facc.x = __fadd_rn(facc.x, __uint_as_float(v[u].x));
                facc.y = __fadd_rn(facc.y, __uint_as_float(v[u].y));
                facc.z = __fadd_rn(facc.z, __uint_as_float(v[u].z));
                facc.w = __fadd_rn(facc.w, __uint_as_float(v[u].w));
            }
        }
    }
    if (bad && !(flags & kDescSharded))
        atomicOr(err_flag, 1u);

    uint4 o;
    if (is_float) {
        o = make_uint4(__float_as_uint(facc.x), __float_as_uint(facc.y), __float_as_uint(facc.z),
                       __float_as_uint(facc.w));
    } else if (flags & kDescRaw) {
        o = acc;
    } else {  // post_process, emb_host.h:210
        o = make_uint4(__float_as_uint(dequant(acc.x)), __float_as_uint(dequant(acc.y)),
                       __float_as_uint(dequant(acc.z)), __float_as_uint(dequant(acc.w)));
    }

    const uint32_t b = b0 + bag;
    uint32_t *base;
    uint32_t row;
    const uint32_t bpr = d->bags_per_rank;
    if (bpr) {  // pooled row goes straight to the slab of the rank that owns sample b
        uint32_t rk = b / bpr;
        base = reinterpret_cast<uint32_t *>(d->out_peer[rk]);
        row = b - rk * bpr;
    } else {
        base = reinterpret_cast<uint32_t *>(d->out);
        row = b;
    }
    uint32_t *dst = base + (size_t)row * d->out_stride + lane * 4;
    if (flags & kDescVecStore) {
        stg_stream16(dst, o);
    } else {
        dst[0] = o.x;
        dst[1] = o.y;
        dst[2] = o.z;
        dst[3] = o.w;
    }
}

/* ---- short-bag kernel -----------------------------------------------------------------------------
 *
 * bag_sum_short_kernel: launches whose bags hold about ONE index each -- the reference's Kaggle shape (README.md:14,
 * dlrm_s_pytorch.py:444: one categorical value per feature and sample), where pooling degenerates to "fetch row
 * indices[b], convert, store".  The stretch kernel's machinery (start markers, heads, the merge of sub-warp partials,
 * shared-memory staging, a barrier per pass) is pure overhead there, and its one-warp CTAs make such a launch ~32 000
 * CTAs of almost no work each (C3: 0.0353 ms, 0.60 of the copy peak; profiles/r3_c3_short_bags.md).
 *
 * Here a warp owns 32 consecutive bags.  Lane l reads the bounds of bag l (one coalesced load of the offsets; the end
 * of a bag is the neighbour's start, taken with a shuffle) and the FIRST index of its bag (coalesced again when every
 * bag holds one index).  The warp then walks its 32 bags in steps of 32/LANES bags -- a sub-warp of LANES lanes per
 * bag, 16 bytes of the row per lane -- with UN steps' row loads issued before any is consumed.  Nothing goes through
 * shared memory and no barrier is needed.  A bag with more than one index is finished by its sub-warp in a plain loop
 * (index order, so fp32 sums are the sequential ones); an empty bag stores zeros.  Arithmetic: emb_dpu_lookup.c:106-116,
 * emb_host.h:207-212, as in every other kernel of this file.  Correct for any bag lengths; chosen by the host only when
 * the launch averages at most kShortBagPool indices per bag.
 */
template <int LANES, int UN, int TPW, int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
bag_sum_short_kernel(const BagDesc *__restrict__ descs, uint32_t n_desc, uint32_t uni, uint32_t *__restrict__ err_flag)
{
    constexpr int GROUPS = 32 / LANES;  // bags a warp handles per step
    constexpr uint32_t FULL = 0xffffffffu;
    const uint32_t q = find_desc(descs, n_desc, uni, blockIdx.x);
    const BagDesc *__restrict__ d = descs + q;
    const uint32_t lane = threadIdx.x & 31u;
    // a warp owns TPW tiles of 32 consecutive bags; the bounds and first indices of ALL its tiles are loaded before
    // any row, so the offsets -> index -> row chain of dependent loads is paid once per warp, not once per tile
    const uint32_t b0 = chunk_first_bag(d, blockIdx.x - d->chunk_begin, kThreads * TPW) + (threadIdx.x >> 5) * (32u * TPW);
    const uint32_t bag_end = d->bag_end;
    if (b0 >= bag_end)
        return;  // whole warp
    const uint32_t B = d->nr_batches;
    const uint32_t ilen = d->indices_len;
    const uint32_t *__restrict__ g_idx = d->indices;
    const uint32_t *__restrict__ g_off = d->offsets;
    const uint32_t row_base = d->row_base;

    // bag bounds, emb_dpu_lookup.c:60-63,109-110: bag 0 starts at 0 whatever offsets[0] is, bag i ends where bag
    // i+1 starts, the last bag ends at indices_len
    uint32_t start[TPW], end[TPW], first[TPW];
#pragma unroll
    for (int t = 0; t < TPW; t++) {
        const uint32_t b = b0 + t * 32u + lane;
        start[t] = 0;
        if (b < bag_end && b != 0)
            start[t] = min(ldg_stream_u32(g_off + b), ilen);
    }
#pragma unroll
    for (int t = 0; t < TPW; t++) {
        const uint32_t b = b0 + t * 32u + lane;
        // the end of a bag is its neighbour's start: lane + 1 of this tile, lane 0 of the next tile, else a load
        uint32_t nxt = __shfl_down_sync(FULL, start[t], 1);
        if (t + 1 < TPW) {
            const uint32_t nt = __shfl_sync(FULL, start[t + 1 < TPW ? t + 1 : t], 0);
            if (lane == 31)
                nxt = nt;
        }
        end[t] = 0;
        if (b < bag_end) {
            const bool have = (lane < 31 || t + 1 < TPW) && b + 1 < bag_end;
            end[t] = (b + 1 >= B) ? ilen : have ? nxt : min(ldg_stream_u32(g_off + b + 1), ilen);
        }
    }
#pragma unroll
    for (int t = 0; t < TPW; t++) {
        first[t] = 0xffffffffu;
        if (end[t] > start[t])
            first[t] = ldg_stream_u32(g_idx + start[t]) - row_base;  // unsigned: rows below a shard wrap to huge values
    }

    const uint32_t flags = d->flags;
    const uint32_t rows = d->rows;
    const bool is_float = flags & kDescFloat;
    const uint32_t grp = lane / LANES, sub = lane % LANES;
    const uint4 *__restrict__ table = reinterpret_cast<const uint4 *>(d->table) + sub;
    const uint32_t bpr = d->bags_per_rank;
    const uint32_t out_stride = d->out_stride;
    bool bad = false;

#pragma unroll
    for (int t = 0; t < TPW; t++) {
        const uint32_t bt = b0 + t * 32u;
        if (bt >= bag_end)
            break;  // whole warp
#pragma unroll 1
        for (int u0 = 0; u0 < LANES; u0 += UN) {
            uint32_t s_[UN], e_[UN];
            uint4 v[UN];
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int j = (u0 + u) * GROUPS + (int)grp;
                s_[u] = __shfl_sync(FULL, start[t], j);
                e_[u] = __shfl_sync(FULL, end[t], j);
                const uint32_t r = __shfl_sync(FULL, first[t], j);
                v[u] = make_uint4(0u, 0u, 0u, 0u);
                if (e_[u] > s_[u]) {
                    if (r < rows)
                        v[u] = ldg_row16<(LANES <= 4)>(table + (size_t)r * LANES);
                    else
                        bad = true;
                }
            }
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const uint32_t bj = bt + (uint32_t)((u0 + u) * GROUPS) + grp;
                if (bj >= bag_end)
                    continue;
                uint4 acc = v[u];
                float4 facc = make_float4(__fadd_rn(0.f, __uint_as_float(v[u].x)), __fadd_rn(0.f, __uint_as_float(v[u].y)),
                                          __fadd_rn(0.f, __uint_as_float(v[u].z)), __fadd_rn(0.f, __uint_as_float(v[u].w)));
                for (uint32_t p = s_[u] + 1; p < e_[u]; p++) {  // the rare longer bag: index order
                    const uint32_t r = __ldg(g_idx + p) - row_base;
                    if (r < rows) {
                        const uint4 x = ldg_row16<(LANES <= 4)>(table + (size_t)r * LANES);
                        acc.x += x.x;  // int32 wrap-around adds (emb_dpu_lookup.c:114)
                        acc.y += x.y;
                        acc.z += x.z;
                        acc.w += x.w;
                        facc.x = __fadd_rn(facc.x, __uint_as_float(x.x));
                        facc.y = __fadd_rn(facc.y, __uint_as_float(x.y));
                        facc.z = __fadd_rn(facc.z, __uint_as_float(x.z));
                        facc.w = __fadd_rn(facc.w, __uint_as_float(x.w));
                    } else {
                        bad = true;
                    }
                }
                uint4 o;
                if (is_float)
                    o = make_uint4(__float_as_uint(facc.x), __float_as_uint(facc.y), __float_as_uint(facc.z),
                                   __float_as_uint(facc.w));
                else if (flags & kDescRaw)
                    o = acc;
                else  // post_process, emb_host.h:210
                    o = make_uint4(__float_as_uint(dequant(acc.x)), __float_as_uint(dequant(acc.y)),
                                   __float_as_uint(dequant(acc.z)), __float_as_uint(dequant(acc.w)));
                uint32_t *base;
                uint32_t row;
                if (bpr) {  // pooled row goes straight to the slab of the rank that owns sample bj
                    const uint32_t rk = bj / bpr;
                    base = reinterpret_cast<uint32_t *>(d->out_peer[rk]);
                    row = bj - rk * bpr;
                } else {
                    base = reinterpret_cast<uint32_t *>(d->out);
                    row = bj;
                }
                uint32_t *dst = base + (size_t)row * out_stride + sub * 4;
                if (flags & kDescVecStore) {
                    stg_stream16(dst, o);
                } else {
                    dst[0] = o.x;
                    dst[1] = o.y;
                    dst[2] = o.z;
                    dst[3] = o.w;
                }
            }
        }
    }
    if (bad && !(flags & kDescSharded))
        atomicOr(err_flag, 1u);
}

/* ---- stretch kernel -----------------------------------------------------------------------------
 *
 * bag_sum_stretch_kernel: load-balanced sum pooling with NO atomics and a branch-free inner loop.
 *
 * The round-1 balanced kernel (pieces of K lookups dealt round-robin to the sub-warps, partial sums flushed into a
 * shared accumulator with atomicAdd at every bag boundary) spent ~400 warp instructions per 64 lookups (ncu: 57 %
 * of the issue slots, 19 % of them branch bookkeeping of the divergent flush, one binary search per piece, 4 ATOMS
 * per flush; profiles/r1_bag_sum_balanced_hotspots.md).  This kernel does the same arithmetic
 * (emb_dpu_lookup.c:106-116, emb_host.h:207-212) in ~100 and replaced it (and the round-1 warp-autonomous
 * kernel) on every workload measured (profiles/r2_stretch_ab_rotated.txt):
 *
 *  - One warp per CTA, BAGS consecutive bags of one table; the bag bounds live in REGISTERS (lane i holds
 *    bound i), so everything that is derived from them -- monotonicity, the window, which bag a position
 *    belongs to -- is a shuffle or a ballot, never a dependent shared-memory search.
 *  - The bags only delimit a window [wlo, whi) of the index stream.  A pass stages up to STAGE lookups of
 *    the window in shared memory: a 4-byte row id and a 1-byte start marker per lookup.  The row id of an index that must
 *    contribute nothing (out of range, another shard's row, padding) is the table's all-zero extra row
 *    (table_alloc), so the gather needs no predicate.  Every non-empty bag then writes "bag + 1" into the
 *    marker of its FIRST lookup: one store per bag, no search per lookup.
 *  - The pass is cut into SUB equal contiguous stretches, one per sub-warp (SUB = 32 / LANES sub-warps of
 *    LANES lanes, a lane owns 16 bytes of every row).  A sub-warp walks its stretch K lookups at a time: the
 *    K row ids and K markers (three shared-memory loads), K independent row loads (LDG.128), then K accumulate
 *    steps.  An accumulate step
 *    is straight-line predicated code: if the lookup carries a marker, the running sum is stored (STS.128)
 *    to where it belongs, the destination moves to the new bag's accumulator and the sum restarts.  The
 *    first such store of a stretch goes to the sub-warp's HEAD slot: the bag its stretch began in started in
 *    an earlier stretch and belongs to that sub-warp.  Every other store targets a bag that started inside
 *    this stretch, which no other sub-warp writes: plain stores, no atomics.
 *  - After the pass each sub-warp adds its head to the accumulator of the bag it continues (one 16-byte
 *    read-modify-write per sub-warp; serialised only when one bag spans several stretches).  A bag cut by
 *    the end of a pass is simply the head of sub-warp 0 in the next pass.
 *  - Epilogue as before: dequantise (post_process, emb_host.h:210) and STG.128 whole output rows.
 *
 * int32 wrap-around sums are exact in any order (bit-exact with the reference).  fp32 tables (IS_FLOAT) are
 * summed in index order inside a stretch and the pieces of a bag are combined in index order too: the result
 * is run-to-run deterministic and differs from the strictly sequential CPU sum (embedding_lookup_idx.cc:33-77)
 * only by the reassociation at the <= SUB-1 cut points of a pass (north_star: <= 1e-5 relative).
 * Bounds that are not monotone inside the CTA's chunk (overlapping bags, which lookup() permits) are pooled bag
 * by bag, as emb_dpu_lookup.c:106-116 does.
 *
 * COMPACT (plans that hold row shards, kDescSharded): a shard owns rows [row_base, row_base + rows) of its table and
 * is handed the table's whole index list.  Its CTAs read the index stream 32 positions at a time, keep the rows of
 * the shard with one ballot per round (slot = running count + popc of the lower lanes) and translate every bag
 * bound inside the round to its place in the KEPT stream the same way; markers, stretches and heads then work on the
 * kept stream exactly as above, so a foreign lookup costs 4 bytes of index traffic and a share of a ballot -- no
 * gather, no accumulate step.  Whole tables in the same launch keep the plain staging.  (Without COMPACT a foreign
 * index gathers the all-zero row behind the table: correct, L1-resident, but a full accumulate step per lookup.)
 * profiles/r2_row_shards.md has the measurements.
 */
/* Programmatic dependent launch (sm_90+).  A kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization
 * may be made resident as soon as every CTA of the kernel in front of it on the stream has executed
 * launch_dependents (or exited); `griddep_wait` then blocks until that kernel has completed and its memory
 * operations are visible.  Without the launch attribute both are no-ops. */
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

/* table + row * ROWB as ONE instruction (IMAD.WIDE.U32 with the 64-bit lane base as the addend) */
template <int ROWB>
__device__ __forceinline__ const uint4 *row_ptr(const char *lane_base, uint32_t row)
{
    uint64_t a;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(a) : "r"(row), "n"(ROWB), "l"(lane_base));
    return reinterpret_cast<const uint4 *>(a);
}
__device__ __forceinline__ uint2 lds_u64(uint32_t smem_addr)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(smem_addr));
    return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t smem_addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_addr));
    return v;
}
__device__ __forceinline__ uint4 lds_u128(uint32_t smem_addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_addr));
    return v;
}
__device__ __forceinline__ void sts_u128(uint32_t smem_addr, uint4 v)
{
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(smem_addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

/* One accumulate step.  mark != 0: the lookup is the first of bag mark-1 -- store the running sum to dst,
 * point dst at that bag's accumulator (acc_base + mark * ROWB; acc_base already holds "- ROWB + lane offset")
 * and restart the sum.  Then add the row.  Predicated, no branch: the sub-warps of a warp meet markers at
 * different lookups. */
template <int ROWB, bool IS_FLOAT>
__device__ __forceinline__ void stretch_step(uint4 &acc, uint32_t &dst, uint32_t mark, uint32_t acc_base, const uint4 &v)
{
    uint32_t keep;  // 0 when a new bag starts at this lookup, else 1
    asm volatile("{\n\t"
                 ".reg .pred p;\n\t"
                 "setp.ne.u32 p, %6, 0;\n\t"
                 "@p st.shared.v4.u32 [%4], {%0,%1,%2,%3};\n\t"
                 "@p mad.lo.u32 %4, %6, %7, %8;\n\t"
                 "selp.u32 %5, 0, 1, p;\n\t"
                 "}"
                 : "+r"(acc.x), "+r"(acc.y), "+r"(acc.z), "+r"(acc.w), "+r"(dst), "=r"(keep)
                 : "r"(mark), "n"(ROWB), "r"(acc_base)
                 : "memory");
    if (IS_FLOAT) {
        const float ax = keep ? __uint_as_float(acc.x) : 0.f, ay = keep ? __uint_as_float(acc.y) : 0.f;
        const float az = keep ? __uint_as_float(acc.z) : 0.f, aw = keep ? __uint_as_float(acc.w) : 0.f;
        acc.x = __float_as_uint(__fadd_rn(ax, __uint_as_float(v.x)));
        acc.y = __float_as_uint(__fadd_rn(ay, __uint_as_float(v.y)));
        acc.z = __float_as_uint(__fadd_rn(az, __uint_as_float(v.z)));
        acc.w = __float_as_uint(__fadd_rn(aw, __uint_as_float(v.w)));
    } else {  // int32 wrap-around adds (emb_dpu_lookup.c:114); "restart" folded into one multiply-add per column
        acc.x = acc.x * keep + v.x;
        acc.y = acc.y * keep + v.y;
        acc.z = acc.z * keep + v.z;
        acc.w = acc.w * keep + v.w;
    }
}

template <bool IS_FLOAT>
__device__ __forceinline__ uint4 add_rows(uint4 a, uint4 b)
{
    if (IS_FLOAT)
        return make_uint4(__float_as_uint(__fadd_rn(__uint_as_float(a.x), __uint_as_float(b.x))),
                          __float_as_uint(__fadd_rn(__uint_as_float(a.y), __uint_as_float(b.y))),
                          __float_as_uint(__fadd_rn(__uint_as_float(a.z), __uint_as_float(b.z))),
                          __float_as_uint(__fadd_rn(__uint_as_float(a.w), __uint_as_float(b.w))));
    return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

template <int LANES, int K, int BAGS, int STAGE, bool IS_FLOAT, int MINB, bool COMPACT = false, bool HOT = false>
__global__ void __launch_bounds__(32, MINB)
bag_sum_stretch_kernel(const BagDesc *__restrict__ descs, uint32_t n_desc, uint32_t uni, uint32_t *__restrict__ err_flag,
                       uint32_t overlap)
{
    // The next kernel on the stream may become resident once the last CTA of this launch has started; this launch
    // itself waits for the kernel in front of it either here or -- `overlap`, EMB_Q_OVERLAP on every query: nothing
    // this launch READS is written by that kernel -- only in front of its result stores, so that its gathers run
    // under the tail of the previous pooling launch and under the exchange's completion wait.
    griddep_launch_dependents();
    if (!overlap)
        griddep_wait();
    constexpr int SUB = 32 / LANES;  // sub-warps
    constexpr int D = LANES * 4;
    constexpr int ROWB = D * 4;      // bytes per row
    constexpr int NB = (BAGS + 31) / 32;  // bound registers per lane
    constexpr int NST = STAGE / 32;       // staging rounds per pass
    static_assert(K == 4 || K == 8, "K is 4 or 8");
    static_assert(STAGE % (SUB * K) == 0 && STAGE % 32 == 0, "a pass splits evenly");
    static_assert(BAGS % SUB == 0 && (BAGS <= 32 || BAGS % 32 == 0), "BAGS");
    // a pass: row ids (4 bytes per lookup) and start markers (1 byte per lookup: bag + 1, or 0)
    static_assert(BAGS <= 254, "a start marker is one byte");
    static_assert(!(COMPACT && HOT), "row shards and hot-row tables are separate instantiations");
    __shared__ __align__(16) uint32_t s_idx[STAGE];
    __shared__ __align__(16) uint8_t s_mark[STAGE];
    __shared__ __align__(16) uint32_t s_acc[BAGS * D];
    __shared__ __align__(16) uint32_t s_head[SUB * D];

    const uint32_t lane = threadIdx.x;
    const uint32_t sub = lane / LANES;
    const uint32_t sl = lane % LANES;
    // Which (query, chunk) is this CTA?  Uniform launches (every query owns `uni` chunks) use a 2-D grid: blockIdx.y is
    // the query, blockIdx.x the chunk, no division.  CTAs are handed out x-fastest, so the launch walks the tables one
    // after the other -- the order everything else was measured against: interleaving the chunks of 2 / 4 / 8
    // neighbouring tables (so that the 64-byte rows of a sample-major result reach L2 as whole lines) costs
    // +11 / +30 / +48 % on C2, because the rows being gathered then come from 2 / 4 / 8 tables at once
    // (profiles/r2_scaling.md).
    uint32_t q, chunk;
    if (uni) {
        q = blockIdx.y;
        chunk = blockIdx.x;
    } else {
        q = find_desc(descs, n_desc, 0, blockIdx.x);
        chunk = blockIdx.x - __ldg(&descs[q].chunk_begin);
    }
    const BagDesc *__restrict__ d = descs + q;

    const uint32_t B = d->nr_batches;
    const uint32_t ilen = d->indices_len;
    const uint32_t b0 = chunk_first_bag(d, chunk, BAGS);
    const uint32_t nb = min((uint32_t)BAGS, d->bag_end - b0);
    const uint32_t *__restrict__ g_idx = d->indices;
    const uint32_t *__restrict__ g_off = d->offsets;

    // Bag bounds in registers (emb_dpu_lookup.c:60-63,109-110: bag 0 starts at 0 whatever offsets[0] is, bag i
    // ends where bag i+1 starts, the last bag ends at indices_len): offv[r] = bound of local bag r*32 + lane,
    // whi = bound nb (the end of the CTA's window), loaded by every lane from one address.
    uint32_t offv[NB];
#pragma unroll
    for (int r = 0; r < NB; r++) {
        const uint32_t i = r * 32 + lane, b = b0 + i;
        uint32_t v = 0;
        if (i < nb && b != 0)
            v = ldg_stream_u32(g_off + b);
        offv[r] = v;
    }
    uint32_t whi = ilen;
    if (b0 + nb < B)
        whi = ldg_stream_u32(g_off + b0 + nb);
    // in the shadow of those loads
    const uint32_t acc_s = (uint32_t)__cvta_generic_to_shared(s_acc);
    for (uint32_t i = lane; i < BAGS * D / 4; i += 32)
        sts_u128(acc_s + i * 16, make_uint4(0u, 0u, 0u, 0u));
    whi = min(whi, ilen);
    bool nonmono = false;
#pragma unroll
    for (int r = 0; r < NB; r++) {
        const uint32_t i = r * 32 + lane;
        offv[r] = (i < nb) ? min(offv[r], ilen) : 0xffffffffu;
    }
    // COMPACT (row shards, kDescSharded): the pass machinery below works on the stream of IN-RANGE lookups only.
    // coff[r] = position in that stream where local bag r*32 + lane starts, known once the index stream has been
    // read up to the bag's bound (0xffffffff before: "not reached yet", which every comparison below treats
    // correctly as "beyond everything staged so far"); wend = its total length, known when the window is exhausted.
    // A whole table in a COMPACT launch keeps stream coordinates (coff = offv) and the usual staging.
    const bool compacting = COMPACT && (d->flags & kDescSharded);
    uint32_t coff[COMPACT ? NB : 1];
    uint32_t wend = compacting ? 0xffffffffu : whi;
#pragma unroll
    for (int r = 0; r < (COMPACT ? NB : 1); r++)
        coff[r] = compacting ? 0xffffffffu : offv[r];
    auto bound = [&](int r) -> uint32_t {
        if constexpr (COMPACT)
            return coff[r];
        else
            return offv[r];
    };
    // bound of the bag after local bag r*32 + lane (the window's end for the last one)
    auto next_bound = [&](int r) {
        uint32_t up = __shfl_down_sync(0xffffffffu, bound(r), 1);
        if (r + 1 < NB) {
            const uint32_t first_of_next = __shfl_sync(0xffffffffu, bound(r + 1 < NB ? r + 1 : r), 0);
            if (lane == 31)
                up = first_of_next;
        }
        return (r * 32 + lane + 1 >= nb) ? wend : up;
    };
#pragma unroll
    for (int r = 0; r < NB; r++) {
        uint32_t up;  // shuffles: every lane takes part
        if constexpr (COMPACT) {
            up = __shfl_down_sync(0xffffffffu, offv[r], 1);
            if (r + 1 < NB) {
                const uint32_t first_of_next = __shfl_sync(0xffffffffu, offv[r + 1 < NB ? r + 1 : r], 0);
                if (lane == 31)
                    up = first_of_next;
            }
            if (r * 32 + lane + 1 >= nb)
                up = whi;
        } else {
            up = next_bound(r);
        }
        nonmono |= r * 32 + lane < nb && offv[r] > up;
    }
    nonmono = __any_sync(0xffffffffu, nonmono);
    const uint32_t wlo = __shfl_sync(0xffffffffu, offv[0], 0);

    const uint32_t flags = d->flags;
    const uint32_t rows = d->rows;
    const uint32_t row_base = d->row_base;
    const char *__restrict__ table = reinterpret_cast<const char *>(d->table) + sl * 16;
    bool bad = false;
    // HOT (hot-row tables): row ids are translated to slab order while they are staged (remap), and every row load
    // carries an L2 policy picked by "slab row < hot_rows".
    const uint32_t *__restrict__ remap = nullptr;
    uint32_t hot_rows = 0;
    uint64_t pol_hot = 0, pol_cold = 0;
    if constexpr (HOT) {
        remap = d->remap;
        hot_rows = d->hot_rows;
        const uint32_t hm = d->hot_mode;
        const uint64_t pol_n = l2_policy(0), pol_l = l2_policy(1), pol_f = l2_policy(2);
        pol_hot = (hm & kHotKeep) ? pol_l : pol_n;
        pol_cold = (hm & kHotStreamCold) ? pol_f : pol_n;
    }

    if (!nonmono) {
        const uint32_t idx_s = (uint32_t)__cvta_generic_to_shared(s_idx);
        const uint32_t mark_s = (uint32_t)__cvta_generic_to_shared(s_mark);
        const uint32_t head_s = (uint32_t)__cvta_generic_to_shared(s_head) + sub * ROWB + sl * 16;
        const uint32_t acc_base = acc_s + sl * 16 - ROWB;  // + mark * ROWB = this lane's 16 bytes of bag mark-1
        // base = first position of the pass, in the coordinates of the bounds (index stream; COMPACT: stream of
        // in-range lookups); src = next position of the index stream to read (COMPACT only)
        uint32_t src = wlo;
        for (uint32_t base = compacting ? 0u : wlo;;) {
            uint32_t n, per;
            uint32_t hb = 0;
            auto head_bags = [&]() {
                // the bag each stretch begins in (= the bag its head belongs to): the last bag whose bound is <= the
                // stretch's first position, counted with one ballot per bound register
#pragma unroll
                for (int s = 0; s < SUB; s++) {
                    const uint32_t pos = base + s * per;
                    uint32_t cnt = 0;
#pragma unroll
                    for (int r = 0; r < NB; r++)
                        cnt += __popc(__ballot_sync(0xffffffffu, bound(r) <= pos));  // bounds past nb are 0xffffffff
                    if ((uint32_t)s == sub)
                        hb = cnt - 1;  // bound 0 <= base, so cnt >= 1
                }
            };
            if (!compacting) {
                if (base >= whi)
                    break;
                n = min((uint32_t)STAGE, whi - base);
                per = K * ((n + SUB * K - 1) / (SUB * K));  // lookups per sub-warp, a multiple of K
                // ---- stage: all index loads of the pass go out back to back
                uint32_t raw[NST];
#pragma unroll
                for (int j = 0; j < NST; j++) {
                    const uint32_t i = j * 32 + lane;
                    raw[j] = 0;
                    if (i < n)
                        raw[j] = ldg_stream_u32(g_idx + base + i);
                }
                if (lane < STAGE / 16)
                    sts_u128(mark_s + lane * 16, make_uint4(0u, 0u, 0u, 0u));
                head_bags();  // while they fly
                if constexpr (HOT) {
                    if (remap) {  // caller's row id -> slab row; all translations of the pass go out back to back
#pragma unroll
                        for (int j = 0; j < NST; j++) {
                            const uint32_t i = j * 32 + lane;
                            const uint32_t r = raw[j] - row_base;
                            if (i < n && r < rows)
                                raw[j] = __ldg(remap + r) + row_base;
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < NST; j++) {
                    const uint32_t i = j * 32 + lane;
                    uint32_t r = raw[j] - row_base;  // unsigned: rows below the shard wrap high
                    const bool in_range = r < rows;
                    bad |= i < n && !in_range;
                    if (i >= n || !in_range)
                        r = rows;  // the all-zero row behind the table
                    s_idx[i] = r;  // every slot of the pass buffer, padding included
                }
            } else if constexpr (COMPACT) {
                // ---- stage: read the index stream 32 positions at a time (up to 4 loads in flight) and keep the
                // rows of this shard, in order, until the pass buffer is full or the window ends.  A ballot per round
                // gives every kept lookup its slot and every bag bound inside the round its place in the kept stream.
                if (lane < STAGE / 16)
                    sts_u128(mark_s + lane * 16, make_uint4(0u, 0u, 0u, 0u));
                n = 0;
                while (src < whi && n + 32 <= (uint32_t)STAGE) {
                    const uint32_t G = min(4u, ((uint32_t)STAGE - n) / 32);
                    uint32_t raw[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t i = src + j * 32 + lane;
                        raw[j] = 0xffffffffu;
                        if ((uint32_t)j < G && i < whi)
                            raw[j] = ldg_stream_u32(g_idx + i);
                    }
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if ((uint32_t)j >= G)
                            break;
                        const uint32_t p0 = src + j * 32;
                        const uint32_t r = raw[j] - row_base;  // unsigned: rows below the shard wrap high
                        const bool keep = p0 + lane < whi && r < rows;
                        const uint32_t m = __ballot_sync(0xffffffffu, keep);
                        if (keep)
                            s_idx[n + __popc(m & ((1u << lane) - 1u))] = r;
#pragma unroll
                        for (int rr = 0; rr < NB; rr++) {
                            const uint32_t o = offv[rr] - p0;  // bounds past nb are 0xffffffff: never inside a round
                            if (o < 32)
                                coff[rr] = base + n + __popc(m & ((1u << o) - 1u));
                        }
                        n += __popc(m);
                    }
                    src += 32 * G;
                }
                if (src >= whi) {  // window exhausted: the kept stream ends here, and so do the bags not reached
                    wend = base + n;
#pragma unroll
                    for (int rr = 0; rr < NB; rr++)
                        if (rr * 32 + lane < nb && coff[rr] == 0xffffffffu)
                            coff[rr] = wend;
                }
                if (n == 0)
                    break;  // nothing kept: only possible once the window is exhausted
                per = K * ((n + SUB * K - 1) / (SUB * K));
                for (uint32_t i = n + lane; i < SUB * per; i += 32)
                    s_idx[i] = rows;  // padding gathers the all-zero row behind the table
                head_bags();
            }
            const bool active = sub * per < n;  // idle sub-warps (all padding) have an all-zero head
            __syncwarp();
            // ---- start markers: one byte store per non-empty bag that begins inside this pass
#pragma unroll
            for (int r = 0; r < NB; r++) {
                const uint32_t i = r * 32 + lane;
                const uint32_t rel = bound(r) - base;
                const uint32_t up = next_bound(r);
                if (i < nb && bound(r) < up && rel < n)  // rel wraps high for bags that began before the pass
                    s_mark[rel] = (uint8_t)(i + 1);
            }
            __syncwarp();
            // ---- gather: every sub-warp walks its stretch, K rows in flight per lane
            {
                uint32_t dst = head_s;
                uint4 acc = make_uint4(0u, 0u, 0u, 0u);
                const uint32_t c0 = sub * per;  // a multiple of K
                // The loop is rotated: the row ids, markers and K row loads of a step are issued at the BOTTOM of
                // the iteration that consumes the previous step, so the loop's back edge sits between a step's
                // loads and its accumulate steps.  Written as one straight block, ptxas sinks every row load to
                // just above the multiply-add that consumes it (one row in flight instead of K; measured 0.374
                // against 0.238 ms on C2).
                uint32_t mk[K / 4];  // K start markers, one byte each
                uint4 v[K];
                auto load_step = [&](uint32_t c) {
                    uint32_t id[K];
#pragma unroll
                    for (int u4 = 0; u4 < K / 4; u4++) {
                        const uint4 t = lds_u128(idx_s + (c0 + c + 4 * u4) * 4);
                        id[4 * u4 + 0] = t.x;
                        id[4 * u4 + 1] = t.y;
                        id[4 * u4 + 2] = t.z;
                        id[4 * u4 + 3] = t.w;
                    }
                    if (K == 8) {
                        const uint2 m = lds_u64(mark_s + c0 + c);
                        mk[0] = m.x;
                        mk[K / 4 - 1] = m.y;
                    } else {
                        mk[0] = lds_u32(mark_s + c0 + c);
                    }
#pragma unroll
                    for (int u = 0; u < K; u++) {
                        if constexpr (HOT)
                            v[u] = ldg_row16_policy<(ROWB <= 64)>(row_ptr<ROWB>(table, id[u]), id[u] < hot_rows, pol_hot,
                                                                  pol_cold);
                        else
                            v[u] = ldg_row16<(ROWB <= 64)>(row_ptr<ROWB>(table, id[u]));
                    }
                };
                load_step(0);
                for (uint32_t c = K; c < per; c += K) {
#pragma unroll
                    for (int u = 0; u < K; u++)
                        stretch_step<ROWB, IS_FLOAT>(acc, dst, (mk[u / 4] >> (8 * (u % 4))) & 0xffu, acc_base, v[u]);
                    load_step(c);
                }
#pragma unroll
                for (int u = 0; u < K; u++)
                    stretch_step<ROWB, IS_FLOAT>(acc, dst, (mk[u / 4] >> (8 * (u % 4))) & 0xffu, acc_base, v[u]);
                sts_u128(dst, acc);  // the bag the stretch ends in (or the head, if no bag began in the stretch)
            }
            __syncwarp();
            // ---- heads: the first piece of every stretch belongs to a bag that began earlier
            {
                const uint32_t hb_prev = __shfl_up_sync(0xffffffffu, hb, LANES);
                const bool dup = __any_sync(0xffffffffu, active && sub > 0 && hb == hb_prev);
                const uint32_t a_addr = acc_s + hb * ROWB + sl * 16;
                if (!dup) {
                    if (active)
                        sts_u128(a_addr, add_rows<IS_FLOAT>(lds_u128(a_addr), lds_u128(head_s)));
                } else {  // one bag spans several stretches: its pieces are added one after the other, in index order
                    for (int s = 0; s < SUB; s++) {
                        if (active && (uint32_t)s == sub)
                            sts_u128(a_addr, add_rows<IS_FLOAT>(lds_u128(a_addr), lds_u128(head_s)));
                        __syncwarp();
                    }
                }
            }
            __syncwarp();
            if (compacting) {
                base += n;
                if (src >= whi)
                    break;
            } else {
                base += STAGE;
            }
        }
    } else {
        // bags overlap or run backwards: pool each bag on its own, as emb_dpu_lookup.c:106-116 does
        static_assert(2 * BAGS <= STAGE, "the fallback keeps 2 bounds per bag in the pass buffer");
        uint32_t *s_off = s_idx;
#pragma unroll
        for (int r = 0; r < NB; r++) {
            const uint32_t i = r * 32 + lane;
            uint32_t up = __shfl_down_sync(0xffffffffu, offv[r], 1);  // (next_bound speaks COMPACT coordinates)
            if (r + 1 < NB) {
                const uint32_t first_of_next = __shfl_sync(0xffffffffu, offv[r + 1 < NB ? r + 1 : r], 0);
                if (lane == 31)
                    up = first_of_next;
            }
            if (i + 1 >= nb)
                up = whi;
            if (i < nb) {
                s_off[2 * i] = offv[r];
                s_off[2 * i + 1] = up;
            }
        }
        __syncwarp();
        for (uint32_t bag = sub; bag < nb; bag += SUB) {
            const uint32_t start = s_off[2 * bag], end = s_off[2 * bag + 1];
            uint4 acc = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t p = start; p < end; p++) {
                uint32_t id = __ldg(g_idx + p) - row_base;
                if (id < rows) {
                    if constexpr (HOT) {
                        if (remap)
                            id = __ldg(remap + id);
                    }
                    acc = add_rows<IS_FLOAT>(acc, ldg_row16<(ROWB <= 64)>(reinterpret_cast<const uint4 *>(table + (size_t)id * ROWB)));
                } else {
                    bad = true;
                }
            }
            sts_u128(acc_s + bag * ROWB + sl * 16, acc);
        }
        __syncwarp();
    }
    if (bad && !(flags & kDescSharded))
        atomicOr(err_flag, 1u);

    const uint32_t bpr = d->bags_per_rank;
    const uint32_t out_stride = d->out_stride;
    if (overlap)
        griddep_wait();
#pragma unroll 2
    for (int it = 0; it < BAGS / SUB; it++) {
        const uint32_t bag = it * SUB + sub;
        if (bag >= nb)
            break;
        uint4 o = lds_u128(acc_s + bag * ROWB + sl * 16);
        if (!IS_FLOAT && !(flags & kDescRaw))  // post_process, emb_host.h:210
            o = make_uint4(__float_as_uint(dequant(o.x)), __float_as_uint(dequant(o.y)),
                           __float_as_uint(dequant(o.z)), __float_as_uint(dequant(o.w)));
        const uint32_t b = b0 + bag;
        uint32_t *obase;
        uint32_t row;
        if (bpr) {
            uint32_t rk = b / bpr;
            obase = reinterpret_cast<uint32_t *>(d->out_peer[rk]);
            row = b - rk * bpr;
        } else {
            obase = reinterpret_cast<uint32_t *>(d->out);
            row = b;
        }
        uint32_t *dstp = obase + (size_t)row * out_stride + sl * 4;
        if (flags & kDescVecStore) {
            stg_stream16(dstp, o);
        } else {
            dstp[0] = o.x;
            dstp[1] = o.y;
            dstp[2] = o.z;
            dstp[3] = o.w;
        }
    }
}

/* ---- TMA (cp.async.bulk.tensor tile::gather4) variant ------------------------------------------
 *
 * bag_sum_tma_kernel: the row gathers are issued to the TMA unit instead of the LSU.  One
 * `cp.async.bulk.tensor.2d ... tile::gather4` (SASS UTMALDG) fetches FOUR rows, named by four row
 * coordinates, through a 2-D tensor map [rows][D] of the table into shared memory and signals an
 * mbarrier with the byte count; rows never pass through registers or L1, so the number of rows in
 * flight is set by shared memory (8 KB per warp here), not by the L1 miss capacity that bounds the
 * LDG kernels.  Row ids outside the table -- a bad index, a row of another
 * shard, the padding of the last group of four -- are simply out-of-bounds coordinates: the TMA
 * unit zero-fills them, which is exactly "contributes nothing".
 *
 * Every warp is autonomous (no CTA barrier): it owns WB = 4 * (32/LANES) consecutive bags, walks their
 * window of the index stream in stages of NR rows through a ring of STAGES shared-memory buffers,
 * and each sub-warp of LANES lanes keeps the running sums of its 4 bags in registers, reading the
 * rows of a bag from shared memory strictly in index order -- so fp32 tables are summed in the same
 * order as the CPU path (embedding_lookup_idx.cc:33-77) and int32 tables need no atomics.
 */
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t phase)
{
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(phase)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void tma_gather4(void *dst_smem, const void *tmap, uint32_t r0, uint32_t r1, uint32_t r2,
                                            uint32_t r3, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes "
                 "[%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(smem_u32(dst_smem)),
                 "l"(tmap), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(smem_u32(bar))
                 : "memory");
}

constexpr int kTmaStages = 4;
constexpr int kTmaStageBytes = 2048;  // per warp and stage: NR = 2048 / row bytes rows

template <int LANES, int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
bag_sum_tma_kernel(const BagDesc *__restrict__ descs, uint32_t n_desc, uint32_t uni, uint32_t *__restrict__ err_flag)
{
    constexpr int ROWB = LANES * 16;
    constexpr int NR = kTmaStageBytes / ROWB;  // rows per stage: 32 (64-byte rows) .. 4 (512-byte rows)
    constexpr int SUBW = 32 / LANES;           // sub-warps per warp
    constexpr int WB = 4 * SUBW;               // bags per warp: 4 per sub-warp
    constexpr int WARPS = kThreads / 32;
    static_assert(NR % 4 == 0 && NR <= 32, "a stage is a whole number of gather4 groups, one row id per lane");
    extern __shared__ __align__(128) uint8_t tma_smem[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t sub = lane / LANES, sl = lane % LANES;
    uint8_t *buf = tma_smem + (size_t)warp * kTmaStages * kTmaStageBytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(tma_smem + (size_t)WARPS * kTmaStages * kTmaStageBytes) + warp * kTmaStages;
    uint32_t *off = reinterpret_cast<uint32_t *>(tma_smem + (size_t)WARPS * kTmaStages * kTmaStageBytes + WARPS * kTmaStages * 8) +
                    warp * (WB + 1);

    const uint32_t q = find_desc(descs, n_desc, uni, blockIdx.x);
    const BagDesc *__restrict__ d = descs + q;
    const uint32_t B = d->nr_batches;
    const uint32_t b0 = chunk_first_bag(d, blockIdx.x - d->chunk_begin, WARPS * WB) + warp * WB;
    if (b0 >= d->bag_end)
        return;
    const uint32_t nb = min((uint32_t)WB, d->bag_end - b0);
    const uint32_t ilen = d->indices_len;
    const uint32_t *__restrict__ g_idx = d->indices;
    const uint32_t *__restrict__ g_off = d->offsets;
    const uint32_t flags = d->flags;
    const uint32_t rows = d->rows;
    const uint32_t row_base = d->row_base;
    const void *tmap = d->tmap;
    const bool is_float = flags & kDescFloat;

    if (lane == 0) {
        for (int s = 0; s < kTmaStages; s++)
            mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    // bag bounds (emb_dpu_lookup.c:60-63,109-110); entries past nb repeat the end of the last bag
    for (uint32_t i = lane; i <= (uint32_t)WB; i += 32) {
        const uint32_t b = b0 + min(i, nb);
        const uint32_t v = (b == 0) ? 0u : (b >= B ? ilen : ldg_stream_u32(g_off + b));
        off[i] = min(v, ilen);
    }
    __syncwarp();
    bool nonmono = false;
    for (uint32_t i = lane; i < nb; i += 32)
        nonmono |= off[i] > off[i + 1];
    nonmono = __any_sync(0xffffffffu, nonmono);

    uint4 acc[4];
#pragma unroll
    for (int j = 0; j < 4; j++)
        acc[j] = make_uint4(0u, 0u, 0u, 0u);
    bool bad = false;

    if (!nonmono) {
        const uint32_t wlo = off[0], whi = off[nb];
        const uint32_t n_stage = (whi - wlo + NR - 1) / NR;

        auto issue = [&](uint32_t k) {
            const uint32_t s = k % kTmaStages;
            const uint32_t p = wlo + k * NR + lane;
            uint32_t id = 0xffffffffu;  // out of bounds: zero-filled by the TMA unit
            if (lane < (uint32_t)NR && p < whi) {
                id = ldg_stream_u32(g_idx + p) - row_base;  // unsigned: rows below the shard wrap high
                if (id >= rows) {
                    bad = true;
                    id = 0xffffffffu;
                }
            }
            if (lane == 0)
                mbar_expect_tx(&bars[s], kTmaStageBytes);
            const uint32_t r0 = __shfl_sync(0xffffffffu, id, (lane * 4 + 0) & 31u);
            const uint32_t r1 = __shfl_sync(0xffffffffu, id, (lane * 4 + 1) & 31u);
            const uint32_t r2 = __shfl_sync(0xffffffffu, id, (lane * 4 + 2) & 31u);
            const uint32_t r3 = __shfl_sync(0xffffffffu, id, (lane * 4 + 3) & 31u);
            if (lane < (uint32_t)(NR / 4))
                tma_gather4(buf + (size_t)s * kTmaStageBytes + (size_t)lane * 4 * ROWB, tmap, r0, r1, r2, r3, &bars[s]);
        };

        for (uint32_t k = 0; k < (uint32_t)kTmaStages && k < n_stage; k++)
            issue(k);
        for (uint32_t k = 0; k < n_stage; k++) {
            const uint32_t s = k % kTmaStages, phase = (k / kTmaStages) & 1u;
            uint32_t spins = 0;
            while (!mbar_try_wait(&bars[s], phase)) {
                if (++spins > (1u << 24)) {  // the copy never arrived: report, do not hang the GPU
                    atomicOr(err_flag, 4u);
                    return;
                }
            }
            const uint32_t p0 = wlo + k * NR, p1 = min(p0 + (uint32_t)NR, whi);
            const uint8_t *src = buf + (size_t)s * kTmaStageBytes + sl * 16;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint32_t bag = sub + SUBW * j;  // interleaved: neighbouring bags go to different sub-warps
                const uint32_t lo = max(off[bag], p0), hi = min(off[bag + 1], p1);
                if (bag < nb) {
                    for (uint32_t p = lo; p < hi; p++) {
                        const uint4 v = *reinterpret_cast<const uint4 *>(src + (size_t)(p - p0) * ROWB);
                        if (!is_float) {  // int32 wrap-around adds (emb_dpu_lookup.c:114)
                            acc[j].x += v.x;
                            acc[j].y += v.y;
                            acc[j].z += v.z;
                            acc[j].w += v.w;
                        } else {          // fp32, strictly in index order
                            acc[j].x = __float_as_uint(__fadd_rn(__uint_as_float(acc[j].x), __uint_as_float(v.x)));
                            acc[j].y = __float_as_uint(__fadd_rn(__uint_as_float(acc[j].y), __uint_as_float(v.y)));
                            acc[j].z = __float_as_uint(__fadd_rn(__uint_as_float(acc[j].z), __uint_as_float(v.z)));
                            acc[j].w = __float_as_uint(__fadd_rn(__uint_as_float(acc[j].w), __uint_as_float(v.w)));
                        }
                    }
                }
            }
            __syncwarp();  // every lane has read stage s: it may be refilled
            if (k + kTmaStages < n_stage)
                issue(k + kTmaStages);
        }
    } else {
        // bags overlap or run backwards: pool each bag on its own, as emb_dpu_lookup.c:106-116 does
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t bag = sub + SUBW * j;
            if (bag >= nb)
                continue;
            const uint4 *__restrict__ table = reinterpret_cast<const uint4 *>(d->table) + sl;
            for (uint32_t p = off[bag]; p < off[bag + 1]; p++) {
                const uint32_t id = __ldg(g_idx + p) - row_base;
                if (id < rows) {
                    const uint4 v = ldg_row16<(LANES <= 4)>(table + (size_t)id * LANES);
                    if (!is_float) {
                        acc[j].x += v.x;
                        acc[j].y += v.y;
                        acc[j].z += v.z;
                        acc[j].w += v.w;
                    } else {
                        acc[j].x = __float_as_uint(__fadd_rn(__uint_as_float(acc[j].x), __uint_as_float(v.x)));
                        acc[j].y = __float_as_uint(__fadd_rn(__uint_as_float(acc[j].y), __uint_as_float(v.y)));
                        acc[j].z = __float_as_uint(__fadd_rn(__uint_as_float(acc[j].z), __uint_as_float(v.z)));
                        acc[j].w = __float_as_uint(__fadd_rn(__uint_as_float(acc[j].w), __uint_as_float(v.w)));
                    }
                } else {
                    bad = true;
                }
            }
        }
    }
    if (bad && !(flags & kDescSharded))
        atomicOr(err_flag, 1u);

    const uint32_t bpr = d->bags_per_rank;
    const uint32_t out_stride = d->out_stride;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const uint32_t bag = sub + SUBW * j;
        if (bag >= nb)
            continue;
        uint4 o = acc[j];
        if (!is_float && !(flags & kDescRaw))  // post_process, emb_host.h:210
            o = make_uint4(__float_as_uint(dequant(o.x)), __float_as_uint(dequant(o.y)),
                           __float_as_uint(dequant(o.z)), __float_as_uint(dequant(o.w)));
        const uint32_t b = b0 + bag;
        uint32_t *obase;
        uint32_t row;
        if (bpr) {
            const uint32_t rk = b / bpr;
            obase = reinterpret_cast<uint32_t *>(d->out_peer[rk]);
            row = b - rk * bpr;
        } else {
            obase = reinterpret_cast<uint32_t *>(d->out);
            row = b;
        }
        uint32_t *dst = obase + (size_t)row * out_stride + sl * 4;
        if (flags & kDescVecStore) {
            stg_stream16(dst, o);
        } else {
            dst[0] = o.x;
            dst[1] = o.y;
            dst[2] = o.z;
            dst[3] = o.w;
        }
    }
}

/*
 * Any embedding dimension (D not a multiple of 4, or above 128): a warp per bag, lanes
 * stride over the columns with 4-byte loads.  Correctness path, not a tuned one.
 */
__global__ void __launch_bounds__(kThreads)
bag_sum_generic_kernel(const BagDesc *__restrict__ descs, uint32_t n_desc, uint32_t uni, uint32_t dim,
                       uint32_t *__restrict__ err_flag)
{
    constexpr int BAGS = kThreads / 32;
    const uint32_t q = find_desc(descs, n_desc, uni, blockIdx.x);
    const BagDesc *__restrict__ d = descs + q;
    const uint32_t B = d->nr_batches;
    const uint32_t b = chunk_first_bag(d, blockIdx.x - d->chunk_begin, BAGS) + threadIdx.x / 32;
    const uint32_t lane = threadIdx.x % 32;
    if (b >= d->bag_end)
        return;
    const uint32_t ilen = d->indices_len;
    uint32_t start = (b == 0) ? 0u : min(__ldg(d->offsets + b), ilen);
    uint32_t end = (b == B - 1) ? ilen : min(__ldg(d->offsets + b + 1), ilen);
    const uint32_t flags = d->flags;
    const uint32_t *table = reinterpret_cast<const uint32_t *>(d->table);
    const uint32_t bpr = d->bags_per_rank;
    uint32_t *base = reinterpret_cast<uint32_t *>(bpr ? d->out_peer[b / bpr] : d->out);
    uint32_t row = bpr ? b % bpr : b;
    bool bad = false;
    for (uint32_t c = lane; c < dim; c += 32) {
        uint32_t acc = 0;
        float facc = 0.f;
        for (uint32_t p = start; p < end; p++) {
            uint32_t id = __ldg(d->indices + p) - d->row_base;
            if (id < d->rows) {
                uint32_t v = __ldg(table + (size_t)id * dim + c);
                acc += v;
                facc = __fadd_rn(facc, __uint_as_float(v));
            } else {
                bad = true;
            }
        }
        uint32_t o = (flags & kDescFloat) ? __float_as_uint(facc)
                     : (flags & kDescRaw) ? acc
                                          : __float_as_uint(dequant(acc));
        base[(size_t)row * d->out_stride + c] = o;
    }
    if (bad && !(flags & kDescSharded))
        atomicOr(err_flag, 1u);
}

/* populate_mram(): column `col` of a table arrives as nr_rows contiguous int32
 * (emb_host.h:165-173 pushed it to DPU table*NR_COLS+col); here it is scattered into the
 * row-major slab. */
__global__ void scatter_column_kernel(uint32_t *__restrict__ slab, const uint32_t *__restrict__ column,
                                      uint64_t nr_rows, uint32_t dim, uint32_t col)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
    for (; i < nr_rows; i += step)
        slab[i * dim + col] = column[i];
}

/* export_emb (dlrm_dpu_pytorch.py:318-320) on the GPU; the slab is already row-major so
 * the reference's transpose disappears.  dst may alias src. */
/* emb_table_remap: row r of the slab moves to row new_of_old[r] of a new slab (4-byte elements, any dim). */
__global__ void permute_rows_kernel(uint32_t *__restrict__ dst, const uint32_t *__restrict__ src,
                                    const uint32_t *__restrict__ new_of_old, uint64_t nr_rows, uint32_t dim)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t step = (uint64_t)gridDim.x * blockDim.x, n = nr_rows * dim;
    for (; i < n; i += step) {
        const uint64_t r = i / dim;
        dst[(uint64_t)new_of_old[r] * dim + (i - r * dim)] = src[i];
    }
}

/* emb_remap_indices: caller row ids -> slab row ids, in place; ids outside the table are left as they are (and are
 * still reported by the pooling launch). */
__global__ void remap_indices_kernel(uint32_t *__restrict__ idx, const uint32_t *__restrict__ new_of_old,
                                     uint64_t n, uint32_t nr_rows)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
    for (; i < n; i += step) {
        const uint32_t r = idx[i];
        if (r < nr_rows)
            idx[i] = __ldg(new_of_old + r);
    }
}

__global__ void quantise_kernel(int32_t *dst, const float *src, uint64_t n)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
    for (; i < n; i += step)
        dst[i] = quantise(src[i]);
}

/* Row-split tables: combine the per-shard partial rows at the receiving rank. */
struct PartList {
    const uint32_t *p[EMB_MAX_PEERS];
};
__global__ void reduce_partials_kernel(PartList parts, uint32_t n_parts, uint32_t part_stride,
                                       uint32_t *__restrict__ out, uint32_t out_stride, uint32_t dim,
                                       uint64_t nr_bags, int is_raw_i32)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t n = nr_bags * dim;
    const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
    for (; i < n; i += step) {
        uint64_t b = i / dim;
        uint32_t c = (uint32_t)(i - b * dim);
        if (is_raw_i32) {
            uint32_t acc = 0;
            for (uint32_t s = 0; s < n_parts; s++)
                acc += parts.p[s][b * part_stride + c];
            out[b * out_stride + c] = __float_as_uint(dequant(acc));
        } else {
            float acc = 0.f;
            for (uint32_t s = 0; s < n_parts; s++)
                acc = __fadd_rn(acc, __uint_as_float(parts.p[s][b * part_stride + c]));
            out[b * out_stride + c] = __float_as_uint(acc);
        }
    }
}

/*
 * Completion step of the fused exchange, ONE launch per batch: every rank's pooling kernel has
 * stored its pooled rows (and raw partial sums of row-split tables) into the peers' slabs over
 * NVLink; this kernel (stream-ordered after the pooling kernel, so those stores have been
 * performed) tells every peer "my rows of batch `epoch` are in your slab", waits until every peer
 * has said the same, and then adds up the partial sums of the row-split tables that landed here.
 * It replaces an NCCL all-reduce of one int (~20 us at 8 ranks) plus one reduction launch per
 * split table.
 *
 * flags.p[r] = rank r's flag block as mapped in THIS process (EMB_MAX_PEERS uint32 slots, slot q
 * written by rank q).  Epochs only grow, a rank can be at most one epoch ahead of a peer, and the
 * comparison is "slot - epoch >= 0" in wrap-around arithmetic, so one slot per peer is enough.
 * Block 0 signals; every block waits on the LOCAL flag block only (blocks of this grid never wait
 * on each other, so the grid need not be co-resident).  A peer that never arrives trips the
 * timeout: bit 1 of the error word is set and the kernel returns instead of hanging the GPU.
 */
struct FlagList {
    uint32_t *p[EMB_MAX_PEERS];
};
struct ReduceJob {
    const uint32_t *parts[EMB_MAX_PEERS];
    uint32_t *out;
    uint32_t n_parts;
    uint32_t part_stride;
    uint32_t out_stride;
    uint32_t is_raw;
    uint64_t nr_bags;
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t globaltimer_ns()
{
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(256)
exchange_finish_kernel(FlagList flags, uint32_t n_ranks, uint32_t my_rank, uint32_t epoch,
                       const ReduceJob *__restrict__ jobs, uint32_t n_jobs, uint32_t dim,
                       uint32_t *__restrict__ err_flag)
{
    const uint32_t tid = threadIdx.x;
    // resident early (programmatic dependent launch): the launch latency of this kernel and of the one behind it
    // overlaps the pooling launch's tail; the signal still follows the COMPLETION of everything in front
    griddep_launch_dependents();
    griddep_wait();
    if (blockIdx.x == 0 && tid < n_ranks && tid != my_rank)
        st_release_sys(flags.p[tid] + my_rank, epoch);
    if (tid < n_ranks && tid != my_rank) {
        const uint32_t *slot = flags.p[my_rank] + tid;
        const uint64_t t0 = globaltimer_ns();
        while ((int32_t)(ld_acquire_sys(slot) - epoch) < 0) {
            if (globaltimer_ns() - t0 > 10000000000ull) {  // 10 s: a peer died
                atomicOr(err_flag, 2u);
                break;
            }
            __nanosleep(64);
        }
    }
    __syncthreads();
    for (uint32_t j = 0; j < n_jobs; j++) {
        const ReduceJob *job = jobs + j;
        const uint64_t n = job->nr_bags * dim;
        const uint32_t np = job->n_parts, ps = job->part_stride, os = job->out_stride;
        uint32_t *__restrict__ out = job->out;
        if (dim % 4 == 0 && ps % 4 == 0 && os % 4 == 0 && n / 4 < 0xffffffffull) {
            // 16 bytes per thread per part, the parts' loads independent of each other, 32-bit index arithmetic
            const uint32_t q = dim / 4, n4 = (uint32_t)(n / 4), ps4 = ps / 4, os4 = os / 4;
            const uint4 *pp[EMB_MAX_PEERS];
#pragma unroll
            for (int s = 0; s < EMB_MAX_PEERS; s++)
                pp[s] = reinterpret_cast<const uint4 *>(job->parts[s < (int)np ? s : 0]);
            const bool raw = job->is_raw != 0;
            for (uint32_t i = blockIdx.x * blockDim.x + tid; i < n4; i += gridDim.x * blockDim.x) {
                const uint32_t b = i / q, c = i - b * q;
                const size_t src = (size_t)b * ps4 + c;
                uint4 v[EMB_MAX_PEERS];
#pragma unroll
                for (int s = 0; s < EMB_MAX_PEERS; s++)
                    if (s < (int)np)
                        v[s] = ldg_stream16(pp[s] + src);
                uint4 acc = v[0];
#pragma unroll
                for (int s = 1; s < EMB_MAX_PEERS; s++)
                    if (s < (int)np) {
                        if (raw)
                            acc = make_uint4(acc.x + v[s].x, acc.y + v[s].y, acc.z + v[s].z, acc.w + v[s].w);
                        else
                            acc = make_uint4(__float_as_uint(__fadd_rn(__uint_as_float(acc.x), __uint_as_float(v[s].x))),
                                             __float_as_uint(__fadd_rn(__uint_as_float(acc.y), __uint_as_float(v[s].y))),
                                             __float_as_uint(__fadd_rn(__uint_as_float(acc.z), __uint_as_float(v[s].z))),
                                             __float_as_uint(__fadd_rn(__uint_as_float(acc.w), __uint_as_float(v[s].w))));
                    }
                if (raw)
                    acc = make_uint4(__float_as_uint(dequant(acc.x)), __float_as_uint(dequant(acc.y)),
                                     __float_as_uint(dequant(acc.z)), __float_as_uint(dequant(acc.w)));
                stg_stream16(reinterpret_cast<uint4 *>(out) + (size_t)b * os4 + c, acc);
            }
            continue;
        }
        for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + tid; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
            const uint64_t b = i / dim;
            const uint32_t c = (uint32_t)(i - b * dim);
            if (job->is_raw) {
                uint32_t acc = 0;
                for (uint32_t s = 0; s < np; s++)
                    acc += job->parts[s][b * ps + c];
                out[b * os + c] = __float_as_uint(dequant(acc));
            } else {
                float acc = 0.f;
                for (uint32_t s = 0; s < np; s++)
                    acc = __fadd_rn(acc, __uint_as_float(job->parts[s][b * ps + c]));
                out[b * os + c] = __float_as_uint(acc);
            }
        }
    }
}

}  // namespace emb
