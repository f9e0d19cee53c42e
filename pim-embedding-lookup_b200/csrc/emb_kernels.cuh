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
    uint64_t pad_;
};

/* First bag of the chunk CTA `cta` (relative to the query) pools; `per` = bags per CTA. */
__device__ __forceinline__ uint32_t chunk_first_bag(const BagDesc *d, uint32_t cta, uint32_t per)
{
    uint32_t c = cta + d->chunk_rot;
    if (c >= d->n_chunks)
        c -= d->n_chunks;
    return d->bag_first + c * per;
}

/* EMB_LD_MODE (build-time switch; measurements in profiles/r1_sweeps.md):
 *   4  rows: ld.global.nc -- L1 allocates, L2 default priority                      (default)
 *   0  rows: ld.global.nc.L1::no_allocate (the first default: "a row is used once, keep it out of L1")
 *   1  rows: L1::no_allocate + L2 evict_last policy; indices/offsets/outputs: L2 evict_first policy
 *   2  rows: L1::no_allocate + .L2::64B prefetch size      3  + .L2::128B prefetch size
 *   5  rows: L1::no_allocate + L2 evict_normal policy      6  L1::no_allocate + L2 evict_last policy (hoisted)
 * ncu shows every sector of an L1::no_allocate load arriving at L2 as evict_first
 * (lts__t_sectors_srcunit_tex_op_read_evict_first_*), so the rows of the table being walked were the
 * first thing L2 dropped.  Plain ld.global.nc is faster on every workload: C2 0.2443 -> 0.2422 ms,
 * C3 0.0496 -> 0.0456, C4 0.658 -> 0.6088, C5 0.5023 -> 0.4819 (tools/ldmode_ab.sh, same box).
 * The index / offset streams and the output rows stay L1::no_allocate: they are used once. */
#ifndef EMB_LD_MODE
#define EMB_LD_MODE 4
#endif
#ifndef EMB_BF_SEARCH
#define EMB_BF_SEARCH 0  // balanced kernel: 1 = branch-free bag search per piece (experiment; BAGS is a power of two)
#endif
#ifndef EMB_CONTIG
#define EMB_CONTIG 0  // balanced kernel: 1 = contiguous stretch per sub-warp instead of round-robin pieces (experiment)
#endif

__device__ __forceinline__ uint64_t l2_policy_evict_last()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

__device__ __forceinline__ uint4 ldg_row16(const uint4 *p)
{
    uint4 r;
#if EMB_LD_MODE == 1
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p), "l"(l2_policy_evict_last()));
#elif EMB_LD_MODE == 2
    asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
#elif EMB_LD_MODE == 3
    asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
#elif EMB_LD_MODE == 4
    asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
#elif EMB_LD_MODE == 5 || EMB_LD_MODE == 6
    uint64_t pol;  // not volatile: one policy per thread, hoisted out of the loops
#if EMB_LD_MODE == 5
    asm("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
#else
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
#endif
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p), "l"(pol));
#else
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
#endif
    return r;
}

__device__ __forceinline__ uint32_t ldg_stream_u32(const uint32_t *p)
{
    uint32_t r;
#if EMB_LD_MODE == 1
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;"
                 : "=r"(r)
                 : "l"(p), "l"(l2_policy_evict_first()));
#else
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
#endif
    return r;
}

__device__ __forceinline__ void stg_stream16(void *p, uint4 v)
{
#if EMB_LD_MODE == 1
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x),
                 "r"(v.y), "r"(v.z), "r"(v.w), "l"(l2_policy_evict_first())
                 : "memory");
#else
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
#endif
}

/* 4-byte asynchronous global->shared copy (LDGSTS): no register, no stall at issue. */
__device__ __forceinline__ void cp_async_u32(uint32_t *smem_dst, const uint32_t *gmem_src)
{
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.wait_all;" ::: "memory");
}

/* emb_host.h:210  (float)s / pow(10,9): int->float RNE, divide in double, round to float.
 * 1e9 is exact in fp32 and double has >= 2*24+2 bits, so this equals one IEEE fp32 divide. */
__device__ __forceinline__ float dequant(uint32_t s)
{
    return __fdiv_rn(__int2float_rn((int)s), 1.0e9f);
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
                v[u] = ldg_row16(table + (size_t)r[u] * LANES);
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

/*
 * Load-balanced sum-pool kernel for int32 fixed-point tables, D = 4*LANES columns.
 *
 * bag_sum_kernel above gives every bag to one sub-warp, so with ragged bags (pooling 1..20)
 * a warp runs as long as its longest bag and the short bags' lanes idle -- fewer row loads in
 * flight than the memory system needs.  Here the CTA's bags only define a window of the index
 * stream; the window is cut into pieces of K consecutive lookups and the pieces are dealt
 * round-robin to the sub-warps.  A sub-warp issues its K row loads back to back (every lane
 * busy whatever the bag lengths), then walks the K rows in order, adding into registers and
 * flushing to a per-CTA shared-memory accumulator whenever the walk crosses a bag boundary.
 * int32 wrap-around addition (emb_dpu_lookup.c:114) is associative and commutative, so the
 * shared-memory atomics reproduce the reference's sums bit for bit in any order.
 *
 * BAGS bags per CTA (a multiple of kThreads/LANES); the window is streamed through the
 * kIdxStage-entry staging buffer in passes, so bags of any length work.
 * Requires monotone offsets inside the CTA's chunk; a chunk that is not monotone (bags that
 * overlap, which lookup() permits) is pooled bag by bag instead.
 */
template <int LANES, int K, int BAGS, int MINB, int THREADS = kThreads>
__global__ void __launch_bounds__(THREADS, MINB)
bag_sum_balanced_kernel(const BagDesc *__restrict__ descs, uint32_t n_desc, uint32_t uni, uint32_t *__restrict__ err_flag)
{
    constexpr int SUB = THREADS / LANES;  // sub-warps per CTA
    constexpr int D = LANES * 4;
    static_assert(BAGS % SUB == 0, "BAGS must be a multiple of the sub-warp count");
    __shared__ uint32_t s_off[BAGS + 1];
    constexpr int STAGE = kIdxStage * THREADS / kThreads;  // indices staged per pass
    __shared__ uint32_t s_idx[STAGE];
    __shared__ __align__(16) uint32_t s_acc[BAGS * D];

    const uint32_t tid = threadIdx.x;
    const uint32_t q = find_desc(descs, n_desc, uni, blockIdx.x);
    const BagDesc *__restrict__ d = descs + q;

    const uint32_t B = d->nr_batches;
    const uint32_t ilen = d->indices_len;
    const uint32_t b0 = chunk_first_bag(d, blockIdx.x - d->chunk_begin, BAGS);
    const uint32_t nb = min((uint32_t)BAGS, d->bag_end - b0);
    const uint32_t *__restrict__ g_idx = d->indices;
    const uint32_t *__restrict__ g_off = d->offsets;

    // Bag bounds (emb_dpu_lookup.c:60-63,109-110: bag 0 starts at 0 whatever offsets[0] is, bag i ends where
    // bag i+1 starts, the last bag ends at indices_len).  First round: the loads are issued here and looked
    // at further down.  Every thread also loads its right-hand neighbour's bound (the same cache line): monotonicity is checked on the loaded
    // values themselves and one barrier closes both the staging and the check.
    const uint32_t bq = b0 + tid;
    uint32_t raw0 = 0, raw1 = 0;
    if (tid <= nb) {
        if (bq != 0 && bq < B)
            raw0 = ldg_stream_u32(g_off + bq);
        if (tid < nb && bq + 1 < B)
            raw1 = ldg_stream_u32(g_off + bq + 1);
    }
    // in the shadow of those loads
    for (uint32_t i = tid; i < BAGS * D / 4; i += THREADS)
        reinterpret_cast<uint4 *>(s_acc)[i] = make_uint4(0u, 0u, 0u, 0u);

    bool nonmono = false;
    if (tid <= nb) {
        const uint32_t v = min((bq == 0) ? 0u : (bq >= B ? ilen : raw0), ilen);
        if (tid < nb)
            nonmono = v > min((bq + 1 >= B) ? ilen : raw1, ilen);
        s_off[tid] = v;
    }
    for (uint32_t i = tid + THREADS; i <= nb; i += THREADS) {
        const uint32_t b = b0 + i;
        const uint32_t v = min((b >= B) ? ilen : ldg_stream_u32(g_off + b), ilen);
        if (i < nb)
            nonmono |= v > min((b + 1 >= B) ? ilen : ldg_stream_u32(g_off + b + 1), ilen);
        s_off[i] = v;
    }
    const bool ordered = __syncthreads_or(nonmono);

    const uint32_t flags = d->flags;
    const uint32_t rows = d->rows;
    const uint32_t row_base = d->row_base;
    const uint32_t sub = tid / LANES;
    const uint32_t lane = tid % LANES;
    const uint4 *__restrict__ table = reinterpret_cast<const uint4 *>(d->table) + lane;
    bool bad = false;

    if (!ordered) {
        const uint32_t wlo = s_off[0], whi = s_off[nb];
        for (uint32_t base = wlo; base < whi; base += STAGE) {
            const uint32_t n = min((uint32_t)STAGE, whi - base);
            if (base != wlo)
                __syncthreads();  // everyone is done with the previous pass's s_idx
            for (uint32_t i = tid; i < n; i += THREADS)
                s_idx[i] = ldg_stream_u32(g_idx + base + i);
            __syncthreads();
#if EMB_CONTIG
            // Experiment (build switch, default off; tools/piece_loop_ab.sh): every sub-warp pools ONE contiguous
            // stretch of the staged window -- one bag search per sub-warp and pass, the running sum kept in
            // registers across its K-lookup steps, a flush only where a bag ends inside the stretch.
            {
                const uint32_t per = (n + SUB - 1) / SUB;
                const uint32_t cb = min(n, sub * per), ce = min(n, cb + per);
                uint32_t lo = 0, hi = nb;
                const uint32_t pfirst = base + cb;
                while (hi - lo > 1) {  // bag of the stretch's first lookup: last i with s_off[i] <= pfirst
                    uint32_t mid = (lo + hi) >> 1;
                    if (s_off[mid] <= pfirst)
                        lo = mid;
                    else
                        hi = mid;
                }
                uint32_t bag = lo, bag_end = s_off[lo + 1];
                uint4 acc = make_uint4(0u, 0u, 0u, 0u);
                for (uint32_t c = cb; c < ce; c += K) {
                    const uint32_t m = min((uint32_t)K, ce - c);
                    uint32_t r[K];
#pragma unroll
                    for (int u = 0; u < K; u++) {
                        uint32_t id = 0xffffffffu;
                        if ((uint32_t)u < m) {
                            id = s_idx[c + u] - row_base;
                            bad |= id >= rows;
                        }
                        r[u] = id;
                    }
                    uint4 v[K];
#pragma unroll
                    for (int u = 0; u < K; u++) {
                        v[u] = make_uint4(0u, 0u, 0u, 0u);
                        if (r[u] < rows)
                            v[u] = ldg_row16(table + (size_t)r[u] * LANES);
                    }
                    const uint32_t p0 = base + c;
#pragma unroll
                    for (int u = 0; u < K; u++) {
                        if ((uint32_t)u < m) {
                            if (p0 + u >= bag_end) {
                                uint32_t *a = s_acc + bag * D + lane * 4;
                                atomicAdd(a + 0, acc.x);
                                atomicAdd(a + 1, acc.y);
                                atomicAdd(a + 2, acc.z);
                                atomicAdd(a + 3, acc.w);
                                acc = make_uint4(0u, 0u, 0u, 0u);
                                do {
                                    bag++;
                                    bag_end = s_off[bag + 1];
                                } while (p0 + u >= bag_end);
                            }
                            acc.x += v[u].x;
                            acc.y += v[u].y;
                            acc.z += v[u].z;
                            acc.w += v[u].w;
                        }
                    }
                }
                if (cb < ce) {
                    uint32_t *a = s_acc + bag * D + lane * 4;
                    atomicAdd(a + 0, acc.x);
                    atomicAdd(a + 1, acc.y);
                    atomicAdd(a + 2, acc.z);
                    atomicAdd(a + 3, acc.w);
                }
            }
#else
            for (uint32_t c = sub * K; c < n; c += SUB * K) {
                const uint32_t m = min((uint32_t)K, n - c);
                uint32_t r[K];
#pragma unroll
                for (int u = 0; u < K; u++) {
                    uint32_t id = 0xffffffffu;
                    if ((uint32_t)u < m) {
                        id = s_idx[c + u] - row_base;  // unsigned: rows below the shard wrap high
                        bad |= id >= rows;
                    }
                    r[u] = id;
                }
                uint4 v[K];
#pragma unroll
                for (int u = 0; u < K; u++) {
                    v[u] = make_uint4(0u, 0u, 0u, 0u);
                    if (r[u] < rows)
                        v[u] = ldg_row16(table + (size_t)r[u] * LANES);
                }
                // bag of the piece's first lookup: last i with s_off[i] <= p (skips empty bags)
                const uint32_t p0 = base + c;
#if EMB_BF_SEARCH
                // Experiment (build switch, default off): branch-free binary lifting, log2(BAGS) predicated steps
                uint32_t lo = 0;
#pragma unroll
                for (int step = BAGS / 2; step >= 1; step >>= 1) {
                    const uint32_t t = lo + step;
                    if (t < nb && s_off[t] <= p0)
                        lo = t;
                }
#else
                uint32_t lo = 0, hi = nb;
                while (hi - lo > 1) {
                    uint32_t mid = (lo + hi) >> 1;
                    if (s_off[mid] <= p0)
                        lo = mid;
                    else
                        hi = mid;
                }
#endif
                uint32_t bag = lo, bag_end = s_off[lo + 1];
                uint4 acc = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                for (int u = 0; u < K; u++) {
                    if ((uint32_t)u < m) {
                        if (p0 + u >= bag_end) {  // crossed into a later bag: flush the partial sum
                            uint32_t *a = s_acc + bag * D + lane * 4;
                            atomicAdd(a + 0, acc.x);
                            atomicAdd(a + 1, acc.y);
                            atomicAdd(a + 2, acc.z);
                            atomicAdd(a + 3, acc.w);
                            acc = make_uint4(0u, 0u, 0u, 0u);
                            do {
                                bag++;
                                bag_end = s_off[bag + 1];
                            } while (p0 + u >= bag_end);
                        }
                        acc.x += v[u].x;
                        acc.y += v[u].y;
                        acc.z += v[u].z;
                        acc.w += v[u].w;
                    }
                }
                uint32_t *a = s_acc + bag * D + lane * 4;
                atomicAdd(a + 0, acc.x);
                atomicAdd(a + 1, acc.y);
                atomicAdd(a + 2, acc.z);
                atomicAdd(a + 3, acc.w);
            }
#endif
        }
    } else {
        // bags overlap or run backwards: pool each bag on its own, as emb_dpu_lookup.c:106-116 does
        for (uint32_t bag = sub; bag < nb; bag += SUB) {
            const uint32_t start = s_off[bag], end = s_off[bag + 1];
            uint4 acc = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t p = start; p < end; p++) {
                uint32_t id = __ldg(g_idx + p) - row_base;
                if (id < rows) {
                    uint4 v = ldg_row16(table + (size_t)id * LANES);
                    acc.x += v.x;
                    acc.y += v.y;
                    acc.z += v.z;
                    acc.w += v.w;
                } else {
                    bad = true;
                }
            }
            reinterpret_cast<uint4 *>(s_acc)[bag * LANES + lane] = acc;
        }
    }
    if (bad && !(flags & kDescSharded))
        atomicOr(err_flag, 1u);
    __syncthreads();

    const uint32_t bpr = d->bags_per_rank;
    const uint32_t out_stride = d->out_stride;
#pragma unroll
    for (int it = 0; it < BAGS / SUB; it++) {
        const uint32_t bag = it * SUB + sub;
        if (bag >= nb)
            break;
        uint4 o = reinterpret_cast<const uint4 *>(s_acc)[bag * LANES + lane];
        if (!(flags & kDescRaw))  // post_process, emb_host.h:210
            o = make_uint4(__float_as_uint(dequant(o.x)), __float_as_uint(dequant(o.y)),
                           __float_as_uint(dequant(o.z)), __float_as_uint(dequant(o.w)));
        const uint32_t b = b0 + bag;
        uint32_t *base;
        uint32_t row;
        if (bpr) {
            uint32_t rk = b / bpr;
            base = reinterpret_cast<uint32_t *>(d->out_peer[rk]);
            row = b - rk * bpr;
        } else {
            base = reinterpret_cast<uint32_t *>(d->out);
            row = b;
        }
        uint32_t *dst = base + (size_t)row * out_stride + lane * 4;
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

/* cp.async groups: commit what has been issued, wait until at most N groups are still pending. */
__device__ __forceinline__ void cp_async_commit()
{
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait_group()
{
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

/*
 * Pipelined load-balanced sum-pool kernel (variant 5) for int32 fixed-point tables, rows up to 128 bytes.
 *
 * bag_sum_balanced_kernel with one warp per CTA spends about half of a CTA's life in front of its first
 * row load: descriptor -> bag bounds -> indices are three dependent memory round trips, and only then do
 * rows fly.  Here a CTA (still one warp, still 64 registers, K rows in flight per lane) owns NCH
 * consecutive sub-chunks of BAGS bags.  The bounds of ALL its bags are loaded in one round trip, and the
 * index window of sub-chunk j+1 is copied into the other half of a two-buffer shared-memory ring with
 * cp.async (no registers, no stall at issue) while the rows of sub-chunk j are gathered.  The chain is
 * paid once per CTA instead of once per BAGS bags; accumulators and the store epilogue stay per
 * sub-chunk, so shared memory per CTA is about what the balanced kernel uses (L1 capacity is what keeps
 * row loads in flight).
 *
 * Same arithmetic as the balanced kernel: pieces of K lookups dealt round-robin to the sub-warps, int32
 * wrap-around adds flushed into a shared accumulator at bag boundaries (emb_dpu_lookup.c:114; associative
 * and commutative, so bit-exact in any order), dequantise + STG.128 epilogue (emb_host.h:207-212).  A
 * CTA whose bounds are not monotone pools bag by bag like emb_dpu_lookup.c:106-116.
 */
template <int LANES, int K, int BAGS, int NCH, int MINB>
__global__ void __launch_bounds__(32, MINB)
bag_sum_pipe_kernel(const BagDesc *__restrict__ descs, uint32_t n_desc, uint32_t uni, uint32_t *__restrict__ err_flag)
{
    constexpr int SUB = 32 / LANES;  // sub-warps per CTA
    constexpr int D = LANES * 4;
    constexpr int ALLB = BAGS * NCH;  // bags per CTA
    constexpr int SI = BAGS * 16;     // index entries per ring buffer; longer windows take extra passes
    static_assert(BAGS % SUB == 0, "BAGS must be a multiple of the sub-warp count");
    __shared__ uint32_t s_off[ALLB + 1];
    __shared__ uint32_t s_idx[2][SI];
    __shared__ __align__(16) uint32_t s_acc[BAGS * D];

    const uint32_t tid = threadIdx.x;
    const uint32_t q = find_desc(descs, n_desc, uni, blockIdx.x);
    const BagDesc *__restrict__ d = descs + q;

    const uint32_t B = d->nr_batches;
    const uint32_t ilen = d->indices_len;
    const uint32_t b0 = chunk_first_bag(d, blockIdx.x - d->chunk_begin, ALLB);
    const uint32_t nb_all = min((uint32_t)ALLB, d->bag_end - b0);
    const uint32_t *__restrict__ g_idx = d->indices;
    const uint32_t *__restrict__ g_off = d->offsets;

    // bounds of all the CTA's bags, one round trip (emb_dpu_lookup.c:60-63,109-110: bag 0 starts at 0,
    // bag i ends where bag i+1 starts, the last bag ends at indices_len); every lane also loads its
    // right-hand neighbour's bound so that monotonicity is known without a second pass
    bool nonmono = false;
    {
        constexpr int R = (ALLB + 1 + 31) / 32;
        uint32_t raw0[R], raw1[R];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const uint32_t i = tid + 32u * r, b = b0 + i;
            raw0[r] = raw1[r] = 0;
            if (i <= nb_all) {
                if (b != 0 && b < B)
                    raw0[r] = ldg_stream_u32(g_off + b);
                if (i < nb_all && b + 1 < B)
                    raw1[r] = ldg_stream_u32(g_off + b + 1);
            }
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            const uint32_t i = tid + 32u * r, b = b0 + i;
            if (i <= nb_all) {
                const uint32_t v = min((b == 0) ? 0u : (b >= B ? ilen : raw0[r]), ilen);
                if (i < nb_all)
                    nonmono |= v > min((b + 1 >= B) ? ilen : raw1[r], ilen);
                s_off[i] = v;
            }
        }
    }
    const bool ordered = __any_sync(0xffffffffu, nonmono);  // true = NOT monotone (name as in the balanced kernel)
    __syncwarp();

    const uint32_t flags = d->flags;
    const uint32_t rows = d->rows;
    const uint32_t row_base = d->row_base;
    const uint32_t sub = tid / LANES;
    const uint32_t lane = tid % LANES;
    const uint4 *__restrict__ table = reinterpret_cast<const uint4 *>(d->table) + lane;
    const uint32_t bpr = d->bags_per_rank;
    const uint32_t out_stride = d->out_stride;
    const uint32_t nch = (nb_all + BAGS - 1) / BAGS;
    bool bad = false;

    // index window of sub-chunk j -> ring buffer j & 1 (first SI entries), asynchronously
    auto issue = [&](uint32_t j) {
        const uint32_t nbj = min((uint32_t)BAGS, nb_all - j * BAGS);
        const uint32_t lo = s_off[j * BAGS], hi = s_off[j * BAGS + nbj];
        const uint32_t n = min((uint32_t)SI, hi - lo);
        uint32_t *dst = s_idx[j & 1];
        for (uint32_t i = tid; i < n; i += 32)
            cp_async_u32(dst + i, g_idx + lo + i);
        cp_async_commit();
    };
    if (!ordered && nch)
        issue(0);

    // one stretch of n staged lookups (window positions base .. base+n) of sub-chunk bags off[0..nbj]
    auto pool = [&](const uint32_t *__restrict__ idx, uint32_t n, uint32_t base, const uint32_t *off, uint32_t nbj) {
        for (uint32_t c = sub * K; c < n; c += SUB * K) {
            const uint32_t m = min((uint32_t)K, n - c);
            uint32_t r[K];
#pragma unroll
            for (int u = 0; u < K; u++) {
                uint32_t id = 0xffffffffu;
                if ((uint32_t)u < m) {
                    id = idx[c + u] - row_base;  // unsigned: rows below the shard wrap high
                    bad |= id >= rows;
                }
                r[u] = id;
            }
            uint4 v[K];
#pragma unroll
            for (int u = 0; u < K; u++) {
                v[u] = make_uint4(0u, 0u, 0u, 0u);
                if (r[u] < rows)
                    v[u] = ldg_row16(table + (size_t)r[u] * LANES);
            }
            // bag of the piece's first lookup: last i with off[i] <= p (skips empty bags)
            const uint32_t p0 = base + c;
            uint32_t lo = 0, hi = nbj;
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (off[mid] <= p0)
                    lo = mid;
                else
                    hi = mid;
            }
            uint32_t bag = lo, bag_end = off[lo + 1];
            uint4 acc = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
            for (int u = 0; u < K; u++) {
                if ((uint32_t)u < m) {
                    if (p0 + u >= bag_end) {  // crossed into a later bag: flush the partial sum
                        uint32_t *a = s_acc + bag * D + lane * 4;
                        atomicAdd(a + 0, acc.x);
                        atomicAdd(a + 1, acc.y);
                        atomicAdd(a + 2, acc.z);
                        atomicAdd(a + 3, acc.w);
                        acc = make_uint4(0u, 0u, 0u, 0u);
                        do {
                            bag++;
                            bag_end = off[bag + 1];
                        } while (p0 + u >= bag_end);
                    }
                    acc.x += v[u].x;
                    acc.y += v[u].y;
                    acc.z += v[u].z;
                    acc.w += v[u].w;
                }
            }
            uint32_t *a = s_acc + bag * D + lane * 4;
            atomicAdd(a + 0, acc.x);
            atomicAdd(a + 1, acc.y);
            atomicAdd(a + 2, acc.z);
            atomicAdd(a + 3, acc.w);
        }
    };

    for (uint32_t j = 0; j < nch; j++) {
        const uint32_t nbj = min((uint32_t)BAGS, nb_all - j * BAGS);
        const uint32_t *off = s_off + j * BAGS;
        for (uint32_t i = tid; i < BAGS * D / 4; i += 32)
            reinterpret_cast<uint4 *>(s_acc)[i] = make_uint4(0u, 0u, 0u, 0u);
        if (!ordered) {
            const uint32_t wlo = off[0], whi = off[nbj];
            if (j + 1 < nch) {  // next sub-chunk's indices fly while this one's rows are gathered
                issue(j + 1);
                cp_async_wait_group<1>();
            } else {
                cp_async_wait_group<0>();
            }
            __syncwarp();  // every lane's copies (and the zeroed accumulator) are visible to the warp
            uint32_t *buf = s_idx[j & 1];
            pool(buf, min((uint32_t)SI, whi - wlo), wlo, off, nbj);
            for (uint32_t base = wlo + SI; base < whi; base += SI) {  // windows longer than a ring buffer
                const uint32_t n = min((uint32_t)SI, whi - base);
                __syncwarp();
                for (uint32_t i = tid; i < n; i += 32)
                    buf[i] = ldg_stream_u32(g_idx + base + i);
                __syncwarp();
                pool(buf, n, base, off, nbj);
            }
        } else {
            // bags overlap or run backwards: pool each bag on its own, as emb_dpu_lookup.c:106-116 does
            __syncwarp();
            for (uint32_t bag = sub; bag < nbj; bag += SUB) {
                const uint32_t start = off[bag], end = off[bag + 1];
                uint4 acc = make_uint4(0u, 0u, 0u, 0u);
                for (uint32_t p = start; p < end; p++) {
                    const uint32_t id = __ldg(g_idx + p) - row_base;
                    if (id < rows) {
                        const uint4 v = ldg_row16(table + (size_t)id * LANES);
                        acc.x += v.x;
                        acc.y += v.y;
                        acc.z += v.z;
                        acc.w += v.w;
                    } else {
                        bad = true;
                    }
                }
                reinterpret_cast<uint4 *>(s_acc)[bag * LANES + lane] = acc;
            }
        }
        __syncwarp();

        // post_process (emb_host.h:207-212) for the sub-chunk's bags
#pragma unroll
        for (int it = 0; it < BAGS / SUB; it++) {
            const uint32_t bag = it * SUB + sub;
            if (bag >= nbj)
                break;
            uint4 o = reinterpret_cast<const uint4 *>(s_acc)[bag * LANES + lane];
            if (!(flags & kDescRaw))
                o = make_uint4(__float_as_uint(dequant(o.x)), __float_as_uint(dequant(o.y)),
                               __float_as_uint(dequant(o.z)), __float_as_uint(dequant(o.w)));
            const uint32_t b = b0 + j * BAGS + bag;
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
            uint32_t *dst = obase + (size_t)row * out_stride + lane * 4;
            if (flags & kDescVecStore) {
                stg_stream16(dst, o);
            } else {
                dst[0] = o.x;
                dst[1] = o.y;
                dst[2] = o.z;
                dst[3] = o.w;
            }
        }
        __syncwarp();  // accumulator and ring buffer j & 1 are free again
    }
    if (bad && !(flags & kDescSharded))
        atomicOr(err_flag, 1u);
}

/*
 * Warp-autonomous load-balanced sum-pool kernel for int32 fixed-point tables with short rows
 * (D = 4*LANES <= 32 columns, i.e. rows of 16..128 bytes).
 *
 * Same idea as bag_sum_balanced_kernel -- the bags only delimit a window of the index stream
 * and the lookups of the window are dealt evenly to the sub-warps -- but the unit is a WARP, not
 * a CTA: a warp owns 32 consecutive bags (one offset per lane), stages its own index window in
 * its own slice of shared memory and never meets a CTA barrier, so the 8 warps of a CTA drift
 * apart and some are always gathering while others stage or store.  While staging, every lane
 * also finds the bag of "its" lookup (binary search over the warp's 33 offsets, lane-parallel) and
 * stores it beside the index; the sub-warps then walk CONTIGUOUS stretches of the window (K rows
 * in flight, 16-byte shared-memory reads of K indices and K bag ids at a time), keep the running
 * sum in registers across a whole bag and flush to the warp's shared accumulator only when the
 * bag id changes -- about one flush per bag instead of one per K lookups.  int32 wrap-around
 * addition is associative and commutative (emb_dpu_lookup.c:114), so the result is bit-exact.
 */
template <int LANES, int K, int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
bag_sum_warp_kernel(const BagDesc *__restrict__ descs, uint32_t n_desc, uint32_t uni, uint32_t *__restrict__ err_flag)
{
    constexpr int SUB = 32 / LANES;  // sub-warps per warp
    constexpr int D = LANES * 4;
    constexpr int WB = 32;           // bags per warp
    constexpr int WARPS = kThreads / 32;
    constexpr int STAGE = LANES >= 8 ? 256 : (384 % (SUB * K) == 0) ? 384 : 512;  // lookups staged per warp and pass (48 KB static smem)
    static_assert(K == 4 || K == 8, "K is 4 or 8");
    static_assert(STAGE % (SUB * K) == 0 && STAGE % 32 == 0, "STAGE must split evenly");
    __shared__ uint32_t s_off[WARPS][WB + 1];
    __shared__ __align__(16) uint32_t s_idx[WARPS][STAGE];
    __shared__ __align__(16) uint8_t s_bag[WARPS][STAGE];
    __shared__ __align__(16) uint32_t s_acc[WARPS][WB * D];

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t sub = lane / LANES;
    const uint32_t sl = lane % LANES;
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
    uint32_t *off = s_off[warp];
    uint32_t *acc_s = s_acc[warp];

    // bag bounds (emb_dpu_lookup.c:60-63,109-110), one per lane; lane 0 also fetches the end of bag 31
    {
        uint32_t b = b0 + lane;
        uint32_t v = (b == 0) ? 0u : (b >= B ? ilen : ldg_stream_u32(g_off + b));
        off[lane] = min(v, ilen);
        if (lane == 0) {
            b = b0 + WB;
            off[WB] = (b >= B) ? ilen : min(ldg_stream_u32(g_off + b), ilen);
        }
    }
#pragma unroll
    for (int i = 0; i < WB * D / 4 / 32; i++)
        reinterpret_cast<uint4 *>(acc_s)[i * 32 + lane] = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
    const bool nonmono = __any_sync(0xffffffffu, lane < nb && off[lane] > off[lane + 1]);

    const uint32_t flags = d->flags;
    const uint32_t rows = d->rows;
    const uint32_t row_base = d->row_base;
    const uint4 *__restrict__ table = reinterpret_cast<const uint4 *>(d->table) + sl;
    bool bad = false;

    if (!nonmono) {
        const uint32_t wlo = off[0], whi = off[nb];
        for (uint32_t base = wlo; base < whi; base += STAGE) {
            const uint32_t n = min((uint32_t)STAGE, whi - base);
            const uint32_t per = ((n + SUB * K - 1) / (SUB * K)) * K;  // lookups per sub-warp, multiple of K
            const uint32_t n_pad = per * SUB;
            if (base != wlo)
                __syncwarp();  // the previous pass's stretches have been walked
            // all of the pass's index loads go out back to back as asynchronous global->shared copies
            for (uint32_t i = lane; i < n; i += 32)
                cp_async_u32(&s_idx[warp][i], g_idx + base + i);
            cp_async_wait_all();
            __syncwarp();
#pragma unroll 4
            for (uint32_t i = lane; i < n_pad; i += 32) {
                uint32_t r = 0xffffffffu;
                const uint32_t p = min(base + i, whi - 1);
                if (i < n) {
                    r = s_idx[warp][i] - row_base;  // unsigned: rows below the shard wrap high
                    if (r >= rows) {
                        bad = true;
                        r = 0xffffffffu;
                    }
                }
                // bag of position p: last j < nb with off[j] <= p (empty bags share a start; the last wins)
                uint32_t lo = 0, hi = nb;
#pragma unroll
                for (int it = 0; it < 5; it++) {
                    const uint32_t mid = (lo + hi) >> 1;
                    const bool act = hi - lo > 1;
                    const bool right = off[mid] <= p;
                    lo = (act && right) ? mid : lo;
                    hi = (act && !right) ? mid : hi;
                }
                s_idx[warp][i] = r;
                s_bag[warp][i] = (uint8_t)lo;
            }
            __syncwarp();
            const uint32_t c0 = sub * per;
            if (c0 < n) {
                uint32_t cur = s_bag[warp][c0];
                uint4 acc = make_uint4(0u, 0u, 0u, 0u);
                for (uint32_t c = c0; c < c0 + per; c += K) {
                    uint32_t r[K], bg[K];
#pragma unroll
                    for (int u4 = 0; u4 < K / 4; u4++) {
                        const uint4 rr = *reinterpret_cast<const uint4 *>(&s_idx[warp][c + 4 * u4]);
                        const uint32_t bb = *reinterpret_cast<const uint32_t *>(&s_bag[warp][c + 4 * u4]);
                        r[4 * u4 + 0] = rr.x;
                        r[4 * u4 + 1] = rr.y;
                        r[4 * u4 + 2] = rr.z;
                        r[4 * u4 + 3] = rr.w;
                        bg[4 * u4 + 0] = bb & 0xffu;
                        bg[4 * u4 + 1] = (bb >> 8) & 0xffu;
                        bg[4 * u4 + 2] = (bb >> 16) & 0xffu;
                        bg[4 * u4 + 3] = bb >> 24;
                    }
                    uint4 v[K];
#pragma unroll
                    for (int u = 0; u < K; u++) {
                        v[u] = make_uint4(0u, 0u, 0u, 0u);
                        if (r[u] != 0xffffffffu)
                            v[u] = ldg_row16(table + (size_t)r[u] * LANES);
                    }
#pragma unroll
                    for (int u = 0; u < K; u++) {
                        if (bg[u] != cur) {  // a later bag begins: flush the finished one
                            uint32_t *a = acc_s + cur * D + sl * 4;
                            atomicAdd(a + 0, acc.x);
                            atomicAdd(a + 1, acc.y);
                            atomicAdd(a + 2, acc.z);
                            atomicAdd(a + 3, acc.w);
                            acc = make_uint4(0u, 0u, 0u, 0u);
                            cur = bg[u];
                        }
                        acc.x += v[u].x;
                        acc.y += v[u].y;
                        acc.z += v[u].z;
                        acc.w += v[u].w;
                    }
                }
                uint32_t *a = acc_s + cur * D + sl * 4;
                atomicAdd(a + 0, acc.x);
                atomicAdd(a + 1, acc.y);
                atomicAdd(a + 2, acc.z);
                atomicAdd(a + 3, acc.w);
            }
        }
    } else {
        // bags overlap or run backwards: pool each bag on its own, as emb_dpu_lookup.c:106-116 does
        for (uint32_t bag = sub; bag < nb; bag += SUB) {
            const uint32_t start = off[bag], end = off[bag + 1];
            uint4 acc = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t p = start; p < end; p++) {
                uint32_t id = __ldg(g_idx + p) - row_base;
                if (id < rows) {
                    uint4 v = ldg_row16(table + (size_t)id * LANES);
                    acc.x += v.x;
                    acc.y += v.y;
                    acc.z += v.z;
                    acc.w += v.w;
                } else {
                    bad = true;
                }
            }
            reinterpret_cast<uint4 *>(acc_s)[bag * LANES + sl] = acc;
        }
    }
    if (bad && !(flags & kDescSharded))
        atomicOr(err_flag, 1u);
    __syncwarp();

    const uint32_t bpr = d->bags_per_rank;
    const uint32_t out_stride = d->out_stride;
#pragma unroll
    for (int it = 0; it < WB / SUB; it++) {
        const uint32_t bag = it * SUB + sub;
        if (bag >= nb)
            break;
        uint4 o = reinterpret_cast<const uint4 *>(acc_s)[bag * LANES + sl];
        if (!(flags & kDescRaw))  // post_process, emb_host.h:210
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
                    const uint4 v = ldg_row16(table + (size_t)id * LANES);
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
