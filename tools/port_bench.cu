// port_bench.cu -- does streaming the 12 B/entry matrix streams through the TMA unit (cp.async.bulk into shared
// memory) instead of through the LSU (256-bit ld.global) leave more of the SM's L2 request port to the x gathers?
//
// Persistent warps, each iteration = one "warp step" of the SpMV kernels: 256 random 8-byte gathers from a 32 MB
// (L2-resident) table, 8 per thread, plus -- depending on the mode -- 3 KB of stream data (1 KB "indices" + 2 KB
// "values") for that step:
//   0  gathers only                       1  + streams by 256-bit ld.global.nc (evict_first), one step ahead
//   2  + streams by TMA bulk copies into a warp-private 2-stage shared-memory ring, read back with LDS.128
//   3  as 2 but the ring is not read (the copies alone)
//   4  streams by LDG only, no gathers    5  streams by TMA only, no gathers
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/port_bench tools/port_bench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int STEP = 256;             // entries per warp step
constexpr int WARPS = 8;
constexpr uint32_t STAGE_BYTES = STEP * 12;

template <int MODE>
__global__ void __launch_bounds__(256) port_kernel(const uint32_t *__restrict__ idx, const uint64_t *__restrict__ val,
                                                   const uint64_t *__restrict__ tab, uint32_t tab_mask, uint64_t nsteps,
                                                   uint64_t *__restrict__ out)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char *ring = smem + (size_t)warp * 2 * STAGE_BYTES;
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + (size_t)WARPS * 2 * STAGE_BYTES) + warp * 2;
    constexpr bool TMA = MODE == 2 || MODE == 3 || MODE == 5, LDG = MODE == 1 || MODE == 4, GATHER = MODE <= 3;
    if (TMA && lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[1])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const uint64_t nwarps = (uint64_t)gridDim.x * WARPS, gw = (uint64_t)blockIdx.x * WARPS + warp;
    const uint64_t per = nsteps / nwarps, s0 = gw * per, s1 = s0 + per;
    uint32_t rng = (uint32_t)(gw * 2654435761u) ^ (lane * 40503u + 17u);
    uint64_t acc = 0;
    uint32_t c[8];
    uint64_t v[8];
    auto tma_issue = [&](uint64_t s, int slot) {
        if (lane == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[slot])), "r"(STAGE_BYTES) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(ring + slot * STAGE_BYTES)), "l"(val + s * STEP), "r"(STEP * 8u), "r"(smem_u32(&bar[slot])) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(ring + slot * STAGE_BYTES + STEP * 8)), "l"(idx + s * STEP), "r"(STEP * 4u), "r"(smem_u32(&bar[slot])) : "memory");
        }
    };
    auto ldg_issue = [&](uint64_t s) {
        const uint32_t *ip = idx + s * STEP + lane * 8;
        const uint64_t *vp = val + s * STEP + lane * 8;
        asm volatile("ld.global.nc.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]), "=r"(c[4]), "=r"(c[5]), "=r"(c[6]), "=r"(c[7]) : "l"(ip));
        asm volatile("ld.global.nc.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(vp));
        asm volatile("ld.global.nc.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(v[4]), "=l"(v[5]), "=l"(v[6]), "=l"(v[7]) : "l"(vp + 4));
    };
    if (TMA && s0 < s1) tma_issue(s0, 0);
    if (LDG && s0 < s1) ldg_issue(s0);
    uint32_t it = 0;
    for (uint64_t s = s0; s < s1; s++, it++) {
        const int slot = it & 1;
        uint64_t cur = 0;
        if (LDG) {  // consume the step loaded one iteration ago, then request the next
#pragma unroll
            for (int k = 0; k < 8; k++) cur ^= v[k] + c[k];
            if (s + 1 < s1) ldg_issue(s + 1);
        }
        if (TMA) {
            if (s + 1 < s1) tma_issue(s + 1, slot ^ 1);
        }
        uint64_t g[8];
        if (GATHER) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                rng += 0x9E3779B9u;  // counter; murmur3 finaliser below: lanes must not stride regularly (L2 slice camping)
                uint32_t h = rng;
                h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
                g[k] = __ldg(reinterpret_cast<const unsigned long long *>(tab + (h & tab_mask)));
            }
        }
        if (TMA) {
            const uint32_t parity = (it >> 1) & 1u;
            uint32_t done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(smem_u32(&bar[slot])), "r"(parity) : "memory");
            if (MODE == 2) {
                const uint4 *vs = reinterpret_cast<const uint4 *>(ring + slot * STAGE_BYTES) + lane * 4;
                const uint4 *cs = reinterpret_cast<const uint4 *>(ring + slot * STAGE_BYTES + STEP * 8) + lane * 2;
#pragma unroll
                for (int k = 0; k < 4; k++) { const uint4 q = vs[k]; cur ^= (uint64_t)q.x + q.y + q.z + q.w; }
#pragma unroll
                for (int k = 0; k < 2; k++) { const uint4 q = cs[k]; cur ^= (uint64_t)q.x + q.y + q.z + q.w; }
            }
            __syncwarp();
        }
        if (GATHER) {
#pragma unroll
            for (int k = 0; k < 8; k++) acc += g[k] * (cur | 1);
        } else acc += cur;
    }
    if (acc == 0x123456789ull) out[0] = acc;
}

template <int MODE>
static void run(const char *name, int sms, int occ, const uint32_t *idx, const uint64_t *val, const uint64_t *tab, uint32_t mask,
                uint64_t nsteps, uint64_t *out)
{
    const size_t smem = (size_t)WARPS * 2 * STAGE_BYTES + WARPS * 2 * 8;
    CK(cudaFuncSetAttribute(port_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int grid = sms * occ;
    port_kernel<MODE><<<grid, 256, smem>>>(idx, val, tab, mask, nsteps, out);
    CK(cudaDeviceSynchronize());
    float best = 1e9f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(a);
        port_kernel<MODE><<<grid, 256, smem>>>(idx, val, tab, mask, nsteps, out);
        cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    const uint64_t done = nsteps / ((uint64_t)grid * WARPS) * ((uint64_t)grid * WARPS) * STEP;
    printf("  %-44s occ %d : %7.3f ms  %7.1f G entries/s  (stream %6.1f GB/s)\n", name, occ, best, done / best / 1e6,
           (MODE == 0 ? 0.0 : 12.0) * done / best / 1e6);
}

int main()
{
    int sms = 0;
    CK(cudaSetDevice(0));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const uint64_t n = 1ull << 28, nsteps = n / STEP;
    uint32_t *idx; uint64_t *val, *tab, *out;
    CK(cudaMalloc(&idx, n * 4 + 256));
    CK(cudaMalloc(&val, n * 8 + 256));
    CK(cudaMalloc(&out, 64));
    CK(cudaMemset(idx, 1, n * 4));
    CK(cudaMemset(val, 1, n * 8));
    for (int mb : {32, 64}) {
        const uint32_t entries = (uint32_t)mb * 1048576u / 8u;
        uint32_t mask = 1;
        while (mask * 2 <= entries) mask *= 2;
        mask -= 1;
        CK(cudaMalloc(&tab, (size_t)(mask + 1) * 8));
        CK(cudaMemset(tab, 1, (size_t)(mask + 1) * 8));
        printf("table %u MB\n", (unsigned)(((size_t)mask + 1) * 8 >> 20));
        for (int occ : {2, 3, 4}) {
            run<0>("gathers only", sms, occ, idx, val, tab, mask, nsteps, out);
            run<1>("gathers + LDG.256 streams", sms, occ, idx, val, tab, mask, nsteps, out);
            run<2>("gathers + TMA streams (ring read by LDS.128)", sms, occ, idx, val, tab, mask, nsteps, out);
            run<3>("gathers + TMA streams (ring not read)", sms, occ, idx, val, tab, mask, nsteps, out);
        }
        run<4>("LDG.256 streams only", sms, 4, idx, val, tab, mask, nsteps, out);
        run<5>("TMA streams only", sms, 4, idx, val, tab, mask, nsteps, out);
        CK(cudaFree(tab));
    }
    return 0;
}
