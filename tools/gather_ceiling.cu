// gather_ceiling.cu -- what the B200 memory system gives for random row gathers, with the bag
// structure taken away: every sub-warp of ROW/16 lanes reads rows at pre-generated random ids,
// K loads in flight, XOR-folds them and writes one word per thread at the end.  The GB/s printed
// is row bytes + index bytes over time; it is the ceiling any pooling kernel for that row size
// can approach.  Measurement aid for DESIGN.md (build: nvcc -arch=sm_100a -O3 -o build/gather_ceiling).
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
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

template <int LANES, int K>
__global__ void __launch_bounds__(256) gather(const uint4 *__restrict__ table, const uint32_t *__restrict__ idx,
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

template <int LANES, int K>
void run(const uint4 *table, uint64_t rows, const uint32_t *idx, uint64_t n_idx, uint32_t *out, int ctas_per_sm)
{
    int grid = 148 * ctas_per_sm;
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; i++)
        gather<LANES, K><<<grid, 256>>>(table, idx, n_idx, out);
    CK(cudaEventRecord(a));
    const int reps = 10;
    for (int i = 0; i < reps; i++)
        gather<LANES, K><<<grid, 256>>>(table, idx, n_idx, out);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    ms /= reps;
    double bytes = (double)n_idx * (LANES * 16 + 4);
    printf("{\"row_bytes\": %d, \"K\": %d, \"ctas_per_sm\": %d, \"rows\": %llu, \"ms\": %.4f, \"GBps\": %.1f, "
           "\"Glookups_per_s\": %.2f}\n",
           LANES * 16, K, ctas_per_sm, (unsigned long long)rows, ms, bytes / ms / 1e6, n_idx / ms / 1e6);
    fflush(stdout);
}

int main(int argc, char **argv)
{
    // usage: gather_ceiling [n_idx] -- sweeps region size x row size; K and CTAs/SM at their best values
    const uint64_t n_idx = argc > 1 ? strtoull(argv[1], nullptr, 10) : 17891328ull;
    const uint64_t max_bytes = 16ull << 30;
    void *table;
    CK(cudaMalloc(&table, max_bytes));
    CK(cudaMemset(table, 1, max_bytes));
    uint32_t *out;
    CK(cudaMalloc(&out, 148 * 16 * 256 * 4));
    std::vector<uint32_t> h(n_idx);
    uint32_t *idx;
    CK(cudaMalloc(&idx, n_idx * 4));
    const uint64_t regions[] = {64ull << 20, 128ull << 20, 256ull << 20, 512ull << 20, 1664000000ull, 4ull << 30,
                                16ull << 30};
    const int lanes[] = {4, 16, 32};
    for (uint64_t region : regions)
        for (int li = 0; li < 3; li++) {
            uint64_t rows = region / (lanes[li] * 16);
            uint64_t s = 88172645463325252ull;
            for (uint64_t i = 0; i < n_idx; i++) {
                s ^= s << 13;
                s ^= s >> 7;
                s ^= s << 17;
                h[i] = (uint32_t)(s % rows);
            }
            CK(cudaMemcpy(idx, h.data(), n_idx * 4, cudaMemcpyHostToDevice));
            const uint4 *t = (const uint4 *)table;
            printf("{\"region_MB\": %llu} ", (unsigned long long)(region >> 20));
            if (lanes[li] == 4)
                run<4, 8>(t, rows, idx, n_idx, out, 4);
            else if (lanes[li] == 16)
                run<16, 4>(t, rows, idx, n_idx, out, 6);
            else
                run<32, 4>(t, rows, idx, n_idx, out, 4);
        }
    return 0;
}
