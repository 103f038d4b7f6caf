// gather_bench.cu -- microbenchmark behind DESIGN.md's "gather ceiling": how many random
// 8-byte reads per second can one B200 sustain from an L2-resident table, by access path?
// SpMV on a uniform-random matrix does one such read per stored entry, so this rate bounds
// GNNZ/s independently of HBM bandwidth.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bench gather_bench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void fill_idx(uint32_t *idx, uint64_t n, uint32_t table)
{
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        idx[i] = (uint32_t)(((unsigned __int128)mix(i) * table) >> 64);
}

// MODE 0: ld.global.nc 8B gather, UNROLL independent gathers per thread
// MODE 1: indices only (no gather): the streaming floor
// MODE 2: two 4-byte gathers of adjacent words? no -- 16-byte gather (v2.u64) of the aligned pair
template <int MODE, int UNROLL>
__global__ void __launch_bounds__(256) gather_kernel(const uint32_t *__restrict__ idx, const uint64_t *__restrict__ tab,
                                                     uint64_t n, uint64_t *__restrict__ out)
{
    uint64_t acc = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x * UNROLL;
    for (uint64_t base = (uint64_t)blockIdx.x * blockDim.x * UNROLL + threadIdx.x * UNROLL; base + UNROLL <= n; base += stride) {
        uint32_t c[UNROLL];
        if (UNROLL % 4 == 0) {
#pragma unroll
            for (int k = 0; k < UNROLL / 4; k++) {
                uint4 v = __ldcs(reinterpret_cast<const uint4 *>(idx + base) + k);
                c[4 * k] = v.x; c[4 * k + 1] = v.y; c[4 * k + 2] = v.z; c[4 * k + 3] = v.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < UNROLL; k++) c[k] = __ldcs(idx + base + k);
        }
        if (MODE == 1) {
#pragma unroll
            for (int k = 0; k < UNROLL; k++) acc += c[k];
        } else if (MODE == 0) {
            uint64_t v[UNROLL];
#pragma unroll
            for (int k = 0; k < UNROLL; k++) v[k] = __ldg(reinterpret_cast<const unsigned long long *>(tab + c[k]));
#pragma unroll
            for (int k = 0; k < UNROLL; k++) acc += v[k];
        } else {
            ulonglong2 v[UNROLL];
#pragma unroll
            for (int k = 0; k < UNROLL; k++) v[k] = __ldg(reinterpret_cast<const ulonglong2 *>(tab + (c[k] & ~1u)));
#pragma unroll
            for (int k = 0; k < UNROLL; k++) acc += v[k].x ^ v[k].y;
        }
    }
    if (acc == 0x1234567) out[0] = acc;
}

// SpMV-shaped traffic without any row logic: per entry a 4-byte index and an 8-byte value are
// streamed (256-bit loads, 8 consecutive entries per thread) and one 8-byte element is gathered.
__global__ void __launch_bounds__(256) spmv_like_kernel(const uint32_t *__restrict__ idx, const uint64_t *__restrict__ val,
                                                        const uint64_t *__restrict__ tab, uint64_t n, uint64_t *__restrict__ out)
{
    const uint64_t base = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (base + 8 > n) return;
    uint32_t c[8]; uint64_t v[8];
    asm volatile("ld.global.nc.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]), "=r"(c[4]), "=r"(c[5]), "=r"(c[6]), "=r"(c[7]) : "l"(idx + base));
    asm volatile("ld.global.nc.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(val + base));
    asm volatile("ld.global.nc.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(v[4]), "=l"(v[5]), "=l"(v[6]), "=l"(v[7]) : "l"(val + base + 4));
    uint64_t acc = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) acc += v[k] * __ldg(reinterpret_cast<const unsigned long long *>(tab + c[k]));
    if (acc == 0x1234567) out[0] = acc;
}
static void run_spmv_like(const uint32_t *idx, const uint64_t *val, const uint64_t *tab, uint64_t n, uint64_t *out)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const unsigned grid = (unsigned)(n / 8 / 256);
    spmv_like_kernel<<<grid, 256>>>(idx, val, tab, n, out);
    float best = 1e9f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(a);
        spmv_like_kernel<<<grid, 256>>>(idx, val, tab, n, out);
        cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    printf("  %-28s                    : %7.3f ms  %7.1f G entries/s  (12 B/entry stream %6.1f GB/s)\n", "spmv-like (idx+val+gather)", best,
           n / best / 1e6, 12.0 * n / best / 1e6);
}

template <int MODE, int UNROLL>
static void run(const char *name, const uint32_t *idx, const uint64_t *tab, uint64_t n, uint64_t *out, int sms, int occ)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int grid = sms * occ;
    gather_kernel<MODE, UNROLL><<<grid, 256>>>(idx, tab, n, out);
    float best = 1e9f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(a);
        gather_kernel<MODE, UNROLL><<<grid, 256>>>(idx, tab, n, out);
        cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    printf("  %-28s occ %d unroll %2d : %7.3f ms  %7.1f G gathers/s  (idx stream %6.1f GB/s)\n", name, occ, UNROLL, best,
           n / best / 1e6, 4.0 * n / best / 1e6);
}

int main(int argc, char **argv)
{
    int dev = 0, sms = 0;
    CK(cudaSetDevice(dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const uint64_t n = 1ull << 28;  // 268M gathers, 1 GB of indices
    uint32_t *idx; uint64_t *tab, *out;
    CK(cudaMalloc(&idx, n * 4 + 256));
    CK(cudaMalloc(&tab, (1ull << 27) * 8));  // up to 1 GB table
    CK(cudaMalloc(&out, 64));
    uint64_t *val;
    CK(cudaMalloc(&val, n * 8 + 256));
    CK(cudaMemset(val, 1, n * 8));
    CK(cudaMemset(tab, 1, (1ull << 27) * 8));
    const double mbs[] = {1, 32, 48, 64, 128};
    for (double mb : mbs) {
        const uint32_t table = (uint32_t)(mb * 1048576 / 8);
        fill_idx<<<sms * 8, 256>>>(idx, n, table);
        CK(cudaDeviceSynchronize());
        printf("table %.0f MB (%u entries)\n", mb, table);
        run<1, 8>("indices only", idx, tab, n, out, sms, 8);
        run<0, 4>("ld.nc 8B", idx, tab, n, out, sms, 8);
        run<0, 8>("ld.nc 8B", idx, tab, n, out, sms, 8);
        run<0, 16>("ld.nc 8B", idx, tab, n, out, sms, 8);
        run<0, 8>("ld.nc 8B", idx, tab, n, out, sms, 4);
        run<2, 8>("ld.nc 16B", idx, tab, n, out, sms, 8);
        run_spmv_like(idx, val, tab, n, out);
    }
    return 0;
}
