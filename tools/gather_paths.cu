// gather_paths.cu -- which on-chip path can serve random 8-byte reads of a dense vector fastest?
// The LSU path (ld.global) tops out at ~1 gather / clk / SM (tools/gather_bench.cu: one L1TEX
// wavefront per distinct 128-byte line).  This probes the alternatives that do not go through
// L1TEX, one mode per invocation so a faulting experiment cannot take the others with it:
//   l2stream  : L2-resident streaming read bandwidth (the LTS cap)
//   smem      : random 8-byte reads from a shared-memory table (LDS.64)
//   bulk      : one 16-byte cp.async.bulk (UBLKCP) per gathered element, global -> shared
//   gather4   : cp.async.bulk.tensor.2d tile::gather4, four 16-byte rows per instruction
//   mix       : half of the warps gather through the LSU, half through gather4
//   lsu       : the baseline (ld.global.nc 8 B)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_paths gather_paths.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

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
__global__ void fill_tab(uint64_t *tab, uint64_t n)
{
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) tab[i] = i;
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t *b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity)
{
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(smem_u32(b)), "r"(parity) : "memory");
}

// ---- l2stream ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) l2stream_kernel(const uint4 *__restrict__ buf, uint64_t n16, int reps, uint64_t *out)
{
    uint32_t acc = 0;
    for (int r = 0; r < reps; r++)
        for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) {
            uint4 v;
            asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));
            acc += v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x12345) out[0] = acc;
}

// ---- lsu baseline ----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) lsu_kernel(const uint32_t *__restrict__ idx, const uint64_t *__restrict__ tab, uint64_t n, uint64_t *out)
{
    uint64_t acc = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x * 8;
    for (uint64_t base = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; base + 8 <= n; base += stride) {
        uint4 a = __ldcs(reinterpret_cast<const uint4 *>(idx + base)), b = __ldcs(reinterpret_cast<const uint4 *>(idx + base) + 1);
        uint32_t c[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        uint64_t v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = __ldg(reinterpret_cast<const unsigned long long *>(tab + c[k]));
#pragma unroll
        for (int k = 0; k < 8; k++) acc += v[k];
    }
    if (acc == 0x1234567) out[0] = acc;
}

// ---- smem ----------------------------------------------------------------------------------
// table of TAB_N doubles in shared memory; each thread does random LDS.64 from it
template <int TAB_N>
__global__ void __launch_bounds__(256) smem_kernel(const uint32_t *__restrict__ idx, const uint64_t *__restrict__ tab, uint64_t n, uint64_t *out)
{
    extern __shared__ uint64_t s_tab[];
    for (int i = threadIdx.x; i < TAB_N; i += blockDim.x) s_tab[i] = tab[i];
    __syncthreads();
    uint64_t acc = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x * 8;
    for (uint64_t base = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; base + 8 <= n; base += stride) {
        uint4 a = __ldcs(reinterpret_cast<const uint4 *>(idx + base)), b = __ldcs(reinterpret_cast<const uint4 *>(idx + base) + 1);
        uint32_t c[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 8; k++) acc += s_tab[c[k] & (TAB_N - 1)];
    }
    if (acc == 0x1234567) out[0] = acc;
}

// ---- bulk: 16-byte cp.async.bulk per element ------------------------------------------------
// each warp: batches of 32 lanes x PER copies; one mbarrier per warp
template <int PER>
__global__ void __launch_bounds__(256) bulk_kernel(const uint32_t *__restrict__ idx, const uint64_t *__restrict__ tab, uint64_t n, uint64_t *out)
{
    __shared__ __align__(128) uint64_t s_dst[8][32 * PER * 2];
    __shared__ uint64_t s_bar[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) mbar_init(&s_bar[warp], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    uint64_t acc = 0;
    uint32_t parity = 0;
    const uint64_t wstride = (uint64_t)gridDim.x * 8 * 32 * PER;
    for (uint64_t base = ((uint64_t)blockIdx.x * 8 + warp) * 32 * PER; base + 32 * PER <= n; base += wstride) {
        uint32_t c[PER];
#pragma unroll
        for (int k = 0; k < PER; k++) c[k] = __ldcs(idx + base + k * 32 + lane);
        if (lane == 0) mbar_expect(&s_bar[warp], 32 * PER * 16);
        __syncwarp();
#pragma unroll
        for (int k = 0; k < PER; k++) {
            const uint64_t *src = tab + (c[k] & ~1u);
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 16, [%2];"
                         ::"r"(smem_u32(&s_dst[warp][(k * 32 + lane) * 2])), "l"(src), "r"(smem_u32(&s_bar[warp])) : "memory");
        }
        mbar_wait(&s_bar[warp], parity);
        parity ^= 1;
#pragma unroll
        for (int k = 0; k < PER; k++) acc += s_dst[warp][(k * 32 + lane) * 2 + (c[k] & 1)];
        __syncwarp();
    }
    if (acc == 0x1234567) out[0] = acc;
}

// ---- gather4 -----------------------------------------------------------------------------------
// lane issues one gather4 for 4 of its elements: 4 rows of 16 bytes -> 64 bytes at a 128-byte slot
template <int PER4>  // gather4 instructions per lane per batch
__device__ __forceinline__ uint64_t gather4_batch(const CUtensorMap *tm, const uint32_t *__restrict__ idx, uint64_t base,
                                                  uint64_t *dst /* warp's [PER4*32*16] */, uint64_t *bar, uint32_t parity, int lane)
{
    uint4 c[PER4];
#pragma unroll
    for (int k = 0; k < PER4; k++) c[k] = __ldcs(reinterpret_cast<const uint4 *>(idx + base) + k * 32 + lane);
    if (lane == 0) mbar_expect(bar, 32 * PER4 * 64);
    __syncwarp();
#pragma unroll
    for (int k = 0; k < PER4; k++) {
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
                     ::"r"(smem_u32(dst + (k * 32 + lane) * 16)), "l"(tm), "r"(smem_u32(bar)), "r"(0), "r"((int)(c[k].x >> 1)),
                     "r"((int)(c[k].y >> 1)), "r"((int)(c[k].z >> 1)), "r"((int)(c[k].w >> 1)) : "memory");
    }
    mbar_wait(bar, parity);
    uint64_t acc = 0;
#pragma unroll
    for (int k = 0; k < PER4; k++) {
        const uint64_t *d = dst + (k * 32 + lane) * 16;
        acc += d[0 + (c[k].x & 1)] + d[2 + (c[k].y & 1)] + d[4 + (c[k].z & 1)] + d[6 + (c[k].w & 1)];
    }
    __syncwarp();
    return acc;
}

template <int PER4, bool MIX>
__global__ void __launch_bounds__(256) gather4_kernel(const __grid_constant__ CUtensorMap tm, const uint32_t *__restrict__ idx,
                                                      const uint64_t *__restrict__ tab, uint64_t n, uint64_t *out, int check)
{
    extern __shared__ __align__(128) uint64_t s_raw[];
    uint64_t *s_dst = s_raw;                                   // [8][PER4*32*16]
    uint64_t *s_bar = s_raw + 8 * PER4 * 32 * 16;              // [8]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) mbar_init(&s_bar[warp], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    uint64_t acc = 0;
    uint32_t parity = 0;
    constexpr int WB = 32 * PER4 * 4;  // elements per warp batch
    const uint64_t wstride = (uint64_t)gridDim.x * 8 * WB;
    for (uint64_t base = ((uint64_t)blockIdx.x * 8 + warp) * WB; base + WB <= n; base += wstride) {
        if (MIX && (warp & 1)) {
            // LSU half: same element count through ld.global
#pragma unroll
            for (int k = 0; k < PER4; k++) {
                uint4 c = __ldcs(reinterpret_cast<const uint4 *>(idx + base) + k * 32 + lane);
                uint64_t v0 = __ldg(reinterpret_cast<const unsigned long long *>(tab + c.x)), v1 = __ldg(reinterpret_cast<const unsigned long long *>(tab + c.y));
                uint64_t v2 = __ldg(reinterpret_cast<const unsigned long long *>(tab + c.z)), v3 = __ldg(reinterpret_cast<const unsigned long long *>(tab + c.w));
                acc += v0 + v1 + v2 + v3;
            }
        } else {
            acc += gather4_batch<PER4>(&tm, idx, base, s_dst + warp * PER4 * 32 * 16, &s_bar[warp], parity, lane);
            parity ^= 1;
        }
    }
    if (check) {  // sum of gathered table entries (tab[i] = i) == sum of indices
        atomicAdd((unsigned long long *)out, (unsigned long long)acc);
    } else if (acc == 0x1234567) out[0] = acc;
}

typedef CUresult (*EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <typename F>
static float time_best(F launch, int reps = 5)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    launch();
    CK(cudaDeviceSynchronize());
    float best = 1e9f;
    for (int r = 0; r < reps; r++) {
        cudaEventRecord(a);
        launch();
        cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    return best;
}

int main(int argc, char **argv)
{
    const char *mode = argc > 1 ? argv[1] : "lsu";
    const double mb = argc > 2 ? atof(argv[2]) : 48.0;
    int sms = 0;
    CK(cudaSetDevice(0));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    if (getenv("FETCH")) {  // L2 fetch granularity hint: 32, 64 or 128 bytes
        CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(getenv("FETCH"))));
        size_t g = 0; CK(cudaDeviceGetLimit(&g, cudaLimitMaxL2FetchGranularity));
        printf("L2 fetch granularity limit = %zu\n", g);
    }
    const uint64_t n = 1ull << 28;
    const uint32_t table = (uint32_t)(mb * 1048576 / 8);
    uint32_t *idx; uint64_t *tab, *out;
    CK(cudaMalloc(&idx, n * 4 + 256));
    CK(cudaMalloc(&tab, (size_t)table * 8 + 256));
    CK(cudaMalloc(&out, 64));
    CK(cudaMemset(out, 0, 64));
    fill_idx<<<sms * 8, 256>>>(idx, n, table);
    fill_tab<<<sms * 8, 256>>>(tab, table);
    CK(cudaDeviceSynchronize());

    if (!strcmp(mode, "l2stream")) {
        const uint64_t n16 = (uint64_t)table * 8 / 16;
        const int reps = 64;
        for (int occ : {4, 8}) {
            float ms = time_best([&] { l2stream_kernel<<<sms * occ, 256>>>((const uint4 *)tab, n16, reps, out); });
            printf("l2stream %5.0f MB occ %d : %.3f ms  %.1f GB/s\n", mb, occ, ms, (double)n16 * 16 * reps / ms / 1e6);
        }
    } else if (!strcmp(mode, "lsu")) {
        float ms = time_best([&] { lsu_kernel<<<sms * 8, 256>>>(idx, tab, n, out); });
        printf("lsu      %5.0f MB : %.3f ms  %.1f G gathers/s\n", mb, ms, n / ms / 1e6);
    } else if (!strcmp(mode, "smem")) {
        constexpr int TAB_N = 8192;  // 64 KB
        CK(cudaFuncSetAttribute(smem_kernel<TAB_N>, cudaFuncAttributeMaxDynamicSharedMemorySize, TAB_N * 8));
        for (int occ : {1, 2, 3}) {
            float ms = time_best([&] { smem_kernel<TAB_N><<<sms * occ, 256, TAB_N * 8>>>(idx, tab, n, out); });
            printf("smem 64 KB table, %d CTA/SM : %.3f ms  %.1f G gathers/s (idx stream %.1f GB/s)\n", occ, ms, n / ms / 1e6, 4.0 * n / ms / 1e6);
        }
    } else if (!strcmp(mode, "bulk")) {
        for (int occ : {2, 4, 8}) {
            float ms = time_best([&] { bulk_kernel<4><<<sms * occ, 256>>>(idx, tab, n, out); });
            printf("bulk16   %5.0f MB per 4 occ %d : %.3f ms  %.1f G gathers/s\n", mb, occ, ms, n / ms / 1e6);
        }
        float ms = time_best([&] { bulk_kernel<8><<<sms * 4, 256>>>(idx, tab, n, out); });
        printf("bulk16   %5.0f MB per 8 occ 4 : %.3f ms  %.1f G gathers/s\n", mb, ms, n / ms / 1e6);
    } else if (!strcmp(mode, "gather4") || !strcmp(mode, "mix")) {
        EncodeTiled enc = nullptr;
        cudaDriverEntryPointQueryResult qr;
        CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&enc, cudaEnableDefault, &qr));
        if (!enc) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
        CUtensorMap tm;
        const cuuint64_t dims[2] = {2, table / 2};
        const cuuint64_t strides[1] = {16};
        const cuuint32_t box[2] = {2, 1};
        const cuuint32_t es[2] = {1, 1};
        CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, tab, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled failed: %d\n", (int)r); return 1; }
        const bool mixm = !strcmp(mode, "mix");
        // correctness: sum of gathered == sum of indices
        {
            constexpr int P = 2;
            const size_t sh = (8 * P * 32 * 16 + 8) * 8;
            CK(cudaFuncSetAttribute(gather4_kernel<P, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
            const uint64_t nchk = 1ull << 22;
            gather4_kernel<P, false><<<64, 256, sh>>>(tm, idx, tab, nchk, out, 1);
            CK(cudaDeviceSynchronize());
            uint64_t got = 0; CK(cudaMemcpy(&got, out, 8, cudaMemcpyDeviceToHost));
            uint32_t *h = (uint32_t *)malloc(nchk * 4); CK(cudaMemcpy(h, idx, nchk * 4, cudaMemcpyDeviceToHost));
            uint64_t want = 0; for (uint64_t i = 0; i < nchk; i++) want += h[i];
            printf("gather4 check: %s (got %llu want %llu)\n", got == want ? "ok" : "MISMATCH", (unsigned long long)got, (unsigned long long)want);
            free(h);
        }
#define RUN_G4(P, OCC)                                                                                                     \
    do {                                                                                                                   \
        const size_t sh = (8 * P * 32 * 16 + 8) * 8;                                                                       \
        if (mixm) {                                                                                                        \
            CK(cudaFuncSetAttribute(gather4_kernel<P, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));       \
            float ms = time_best([&] { gather4_kernel<P, true><<<sms * OCC, 256, sh>>>(tm, idx, tab, n, out, 0); });       \
            printf("mix lsu+gather4 %5.0f MB per %d occ %d : %.3f ms  %.1f G gathers/s\n", mb, P, OCC, ms, n / ms / 1e6);     \
        } else {                                                                                                           \
            CK(cudaFuncSetAttribute(gather4_kernel<P, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));      \
            float ms = time_best([&] { gather4_kernel<P, false><<<sms * OCC, 256, sh>>>(tm, idx, tab, n, out, 0); });      \
            printf("gather4  %5.0f MB per %d occ %d : %.3f ms  %.1f G gathers/s\n", mb, P, OCC, ms, n / ms / 1e6);             \
        }                                                                                                                  \
    } while (0)
        RUN_G4(1, 4);
        RUN_G4(2, 2);
        RUN_G4(2, 4);
        RUN_G4(4, 2);
    } else {
        printf("unknown mode %s\n", mode);
        return 1;
    }
    return 0;
}
