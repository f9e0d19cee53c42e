// gather_tma.cu -- LDG.128 vs TMA for random row gathers (the north-star's "TMA gather or plain
// LDG.128, whichever measures faster").  Same experiment as gather_ceiling.cu (no bag structure,
// rows at pre-generated uniform random ids, XOR-folded) with three ways of fetching a row:
//
//   ldg   : ROW/16 lanes per row, K x LDG.128 (ld.global.nc.L1::no_allocate) in flight per lane
//   bulk  : one cp.async.bulk (UBLKCP, 1-D TMA copy of ROW bytes) per row into shared memory,
//           per-warp multi-stage mbarrier pipeline, NR rows per stage
//   g4    : cp.async.bulk.tensor.2d tile::gather4 (UTMALDG): 4 rows per instruction through a 2-D
//           tensor map [rows][ROW/4] int32, same pipeline
//
// The XOR of everything fetched must agree between the modes (printed as "xor").
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o build/gather_tma tools/gather_tma.cu
// usage: gather_tma <mode> <region_MB> <row_bytes> [NR=32] [STAGES=4] [CTAS_PER_SM=3] [BOXH=1] [n_idx]
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e = (x);                                                           \
        if (e != cudaSuccess) {                                                        \
            fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); \
            exit(1);                                                                   \
        }                                                                              \
    } while (0)

__device__ __forceinline__ uint4 ldg16(const uint4 *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

__global__ void fill(uint32_t *t, uint64_t n)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t x = i * 0x9E3779B97F4A7C15ull;
        t[i] = (uint32_t)(x >> 29) ^ (uint32_t)i;
    }
}

template <int LANES, int K>
__global__ void __launch_bounds__(256) gather_ldg(const uint4 *__restrict__ table, const uint32_t *__restrict__ idx,
                                                  uint64_t n_idx, uint32_t *__restrict__ out)
{
    const uint32_t lane = threadIdx.x % LANES;
    const uint64_t sub = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES;
    const uint64_t nsub = (uint64_t)gridDim.x * blockDim.x / LANES;
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (uint64_t p = sub * K; p + K <= n_idx; p += nsub * K) {
        uint32_t r[K];
#pragma unroll
        for (int u = 0; u < K; u++)
            r[u] = __ldg(idx + p + u);
        uint4 v[K];
#pragma unroll
        for (int u = 0; u < K; u++)
            v[u] = ldg16(table + (size_t)r[u] * LANES + lane);
#pragma unroll
        for (int u = 0; u < K; u++) {
            acc.x ^= v[u].x;
            acc.y ^= v[u].y;
            acc.z ^= v[u].z;
            acc.w ^= v[u].w;
        }
    }
    out[(uint64_t)blockIdx.x * blockDim.x + threadIdx.x] = acc.x ^ acc.y ^ acc.z ^ acc.w;
}

// ---- mbarrier / TMA primitives (PTX) -------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t phase)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra D;\n"
        "bra W;\n"
        "D:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(phase)
        : "memory");
}
__device__ __forceinline__ void bulk_copy(void *dst_smem, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_gather4(void *dst_smem, const CUtensorMap *tm, int col, uint4 rows, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes "
                 "[%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(smem_u32(dst_smem)),
                 "l"(tm), "r"(col), "r"(rows.x), "r"(rows.y), "r"(rows.z), "r"(rows.w), "r"(smem_u32(bar))
                 : "memory");
}

// MODE 1 = bulk (1-D copy per row), 2 = gather4.  ROWB = row bytes.  Every warp runs its own pipeline:
// STAGES buffers of NR rows, one mbarrier each; chunk c of the id stream = ids [c*NR, (c+1)*NR).
template <int MODE, int ROWB>
__global__ void __launch_bounds__(256) gather_tma(const uint8_t *__restrict__ table, const __grid_constant__ CUtensorMap tm,
                                                  const uint32_t *__restrict__ idx, uint64_t n_idx,
                                                  uint32_t *__restrict__ out, int NR, int STAGES)
{
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int LPR = ROWB / 16;   // lanes that read one row back from shared memory
    constexpr int RPP = 32 / LPR;    // rows read per warp pass
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t stage_bytes = (uint32_t)NR * ROWB;
    uint8_t *buf = smem + (size_t)warp * STAGES * stage_bytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)8 * STAGES * stage_bytes) + warp * STAGES;
    if (lane == 0) {
        for (int s = 0; s < STAGES; s++)
            mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();
    const uint64_t gw = (uint64_t)blockIdx.x * 8 + warp, nw = (uint64_t)gridDim.x * 8;
    const uint64_t n_chunks = n_idx / NR;

    auto issue = [&](uint64_t c, int s) {
        if (lane == 0)
            mbar_expect_tx(&bars[s], stage_bytes);
        __syncwarp();
        const uint32_t *ids = idx + c * NR;
        uint8_t *dst = buf + (size_t)s * stage_bytes;
        if (MODE == 1) {
            for (int i = lane; i < NR; i += 32)
                bulk_copy(dst + (size_t)i * ROWB, table + (size_t)__ldg(ids + i) * ROWB, ROWB, &bars[s]);
        } else {
            for (int i = lane; i < NR / 4; i += 32) {
                const uint4 r = __ldg(reinterpret_cast<const uint4 *>(ids) + i);
                tma_gather4(dst + (size_t)i * 4 * ROWB, &tm, 0, r, &bars[s]);
            }
        }
    };

    uint64_t c_issue = gw;
    for (int s = 0; s < STAGES; s++, c_issue += nw)
        if (c_issue < n_chunks)
            issue(c_issue, s);
    uint4 acc = make_uint4(0, 0, 0, 0);
    int s = 0;
    uint32_t phase = 0;
    for (uint64_t c = gw; c < n_chunks; c += nw) {
        mbar_wait(&bars[s], phase);
        const uint8_t *src = buf + (size_t)s * stage_bytes;
        for (int r0 = 0; r0 < NR; r0 += RPP) {
            const uint4 v = *reinterpret_cast<const uint4 *>(src + (size_t)(r0 + lane / LPR) * ROWB + (lane % LPR) * 16);
            acc.x ^= v.x;
            acc.y ^= v.y;
            acc.z ^= v.z;
            acc.w ^= v.w;
        }
        __syncwarp();
        if (c_issue < n_chunks)
            issue(c_issue, s);
        c_issue += nw;
        if (++s == STAGES) {
            s = 0;
            phase ^= 1;
        }
    }
    out[(uint64_t)blockIdx.x * blockDim.x + threadIdx.x] = acc.x ^ acc.y ^ acc.z ^ acc.w;
}

typedef CUresult (*EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static uint32_t xor_all(const uint32_t *d_out, size_t n)
{
    std::vector<uint32_t> h(n);
    CK(cudaMemcpy(h.data(), d_out, n * 4, cudaMemcpyDeviceToHost));
    uint32_t x = 0;
    for (uint32_t v : h)
        x ^= v;
    return x;
}

template <int ROWB>
static void run(const char *mode, uint64_t region_mb, int NR, int STAGES, int ctas, int boxh, uint64_t n_idx)
{
    const uint64_t rows = (region_mb << 20) / ROWB;
    uint8_t *table;
    CK(cudaMalloc(&table, rows * ROWB));
    fill<<<148 * 8, 256>>>((uint32_t *)table, rows * ROWB / 4);
    std::vector<uint32_t> h(n_idx);
    uint64_t s = 88172645463325252ull;
    for (uint64_t i = 0; i < n_idx; i++) {
        s ^= s << 13;
        s ^= s >> 7;
        s ^= s << 17;
        h[i] = (uint32_t)(s % rows);
    }
    uint32_t *idx, *out;
    CK(cudaMalloc(&idx, n_idx * 4));
    CK(cudaMemcpy(idx, h.data(), n_idx * 4, cudaMemcpyHostToDevice));
    const int grid = 148 * ctas;
    const size_t n_out = (size_t)grid * 256;
    CK(cudaMalloc(&out, n_out * 4));
    CK(cudaMemset(out, 0, n_out * 4));

    CUtensorMap tm;
    memset(&tm, 0, sizeof tm);
    const bool is_ldg = !strcmp(mode, "ldg"), is_g4 = !strcmp(mode, "g4");
    if (is_g4) {
        EncodeTiled enc = nullptr;
        cudaDriverEntryPointQueryResult qr;
        CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&enc, cudaEnableDefault, &qr));
        cuuint64_t gdim[2] = {ROWB / 4, rows};
        cuuint64_t gstr[1] = {ROWB};
        cuuint32_t box[2] = {ROWB / 4, (cuuint32_t)boxh};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_INT32, 2, table, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            fprintf(stderr, "cuTensorMapEncodeTiled failed: %d\n", (int)r);
            exit(2);
        }
    }
    const size_t smem = (size_t)8 * STAGES * NR * ROWB + 8 * STAGES * 8;
    auto launch = [&]() {
        if (is_ldg) {
            if (ROWB == 64)
                gather_ldg<ROWB / 16, 8><<<grid, 256>>>((const uint4 *)table, idx, n_idx / NR * NR, out);
            else
                gather_ldg<ROWB / 16, 4><<<grid, 256>>>((const uint4 *)table, idx, n_idx / NR * NR, out);
        } else if (is_g4) {
            gather_tma<2, ROWB><<<grid, 256, smem>>>(table, tm, idx, n_idx, out, NR, STAGES);
        } else {
            gather_tma<1, ROWB><<<grid, 256, smem>>>(table, tm, idx, n_idx, out, NR, STAGES);
        }
    };
    if (!is_ldg) {
        CK(cudaFuncSetAttribute(gather_tma<1, ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CK(cudaFuncSetAttribute(gather_tma<2, ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    launch();
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    const uint32_t x = xor_all(out, n_out);
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; i++)
        launch();
    CK(cudaEventRecord(a));
    const int reps = 10;
    for (int i = 0; i < reps; i++)
        launch();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    ms /= reps;
    const double n = (double)(n_idx / NR * NR);
    printf("{\"mode\": \"%s\", \"region_MB\": %llu, \"row_bytes\": %d, \"NR\": %d, \"stages\": %d, \"ctas_per_sm\": %d, "
           "\"boxh\": %d, \"smem_per_cta\": %zu, \"ms\": %.4f, \"GBps\": %.1f, \"Grows_per_s\": %.2f, \"xor\": \"%08x\"}\n",
           mode, (unsigned long long)region_mb, ROWB, NR, STAGES, ctas, boxh, is_ldg ? (size_t)0 : smem, ms,
           n * (ROWB + 4) / ms / 1e6, n / ms / 1e6, x);
}

int main(int argc, char **argv)
{
    if (argc < 4) {
        fprintf(stderr, "usage: gather_tma <ldg|bulk|g4> <region_MB> <row_bytes> [NR] [STAGES] [CTAS_PER_SM] [BOXH] [n_idx]\n");
        return 1;
    }
    const char *mode = argv[1];
    const uint64_t region = strtoull(argv[2], nullptr, 10);
    const int rowb = atoi(argv[3]);
    const int NR = argc > 4 ? atoi(argv[4]) : 32, ST = argc > 5 ? atoi(argv[5]) : 4, ctas = argc > 6 ? atoi(argv[6]) : 3;
    const int boxh = argc > 7 ? atoi(argv[7]) : 1;
    const uint64_t n_idx = argc > 8 ? strtoull(argv[8], nullptr, 10) : 17891328ull;
    if (NR % 32 || NR <= 0 || ST <= 0) {
        fprintf(stderr, "NR must be a positive multiple of 32\n");
        return 1;
    }
    if (rowb == 64)
        run<64>(mode, region, NR, ST, ctas, boxh, n_idx);
    else if (rowb == 256)
        run<256>(mode, region, NR, ST, ctas, boxh, n_idx);
    else if (rowb == 512)
        run<512>(mode, region, NR, ST, ctas, boxh, n_idx);
    else
        return 1;
    return 0;
}
