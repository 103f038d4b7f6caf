// scan.cu -- in-place exclusive prefix sum over u32 (reduce-then-scan, recursive on block sums).
// Device statement of the running sum in software/csr2csc.c:22 ("col_start[i+1] += col_start[i]").
#include "common.cuh"

namespace {

constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 16;  // per thread, as 4 x uint4
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) >= d) v += t;
    }
    return v;
}

// block-wide exclusive scan of one value per thread; returns exclusive prefix, total in *total
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *total)
{
    __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = warp_incl_scan(v);
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < SCAN_THREADS / 32 ? warp_sums[lane] : 0;
        uint32_t wi = warp_incl_scan(w);
        if (lane < SCAN_THREADS / 32) warp_sums[lane] = wi - w;  // exclusive
        if (lane == SCAN_THREADS / 32 - 1) *total = wi;
    }
    __syncthreads();
    uint32_t r = incl - v + warp_sums[warp];
    return r;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(const uint32_t *__restrict__ data,
                                                                  uint64_t n, uint32_t *__restrict__ sums)
{
    __shared__ uint32_t s_total;
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t acc = 0;
    if (base + SCAN_ITEMS <= n) {
        const uint4 *p = reinterpret_cast<const uint4 *>(data + base);
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS / 4; i++) {
            uint4 v = p[i];
            acc += v.x + v.y + v.z + v.w;
        }
    } else {
        for (int i = 0; i < SCAN_ITEMS; i++)
            if (base + i < n) acc += data[base + i];
    }
    block_excl_scan(acc, &s_total);
    if (threadIdx.x == 0) sums[blockIdx.x] = s_total;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(uint32_t *__restrict__ data, uint64_t n,
                                                                 const uint32_t *__restrict__ sums)
{
    __shared__ uint32_t s_total;
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    const bool full = base + SCAN_ITEMS <= n;
    if (full) {
        const uint4 *p = reinterpret_cast<const uint4 *>(data + base);
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS / 4; i++) {
            uint4 q = p[i];
            v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (base + i < n) ? data[base + i] : 0;
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) acc += v[i];
    uint32_t run = block_excl_scan(acc, &s_total) + (sums ? sums[blockIdx.x] : 0u);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        uint32_t t = v[i];
        v[i] = run;
        run += t;
    }
    if (full) {
        uint4 *p = reinterpret_cast<uint4 *>(data + base);
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS / 4; i++) p[i] = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++)
            if (base + i < n) data[base + i] = v[i];
    }
}

}  // namespace

int b2_exclusive_scan_u32(b200spmv_ctx *ctx, uint32_t *data, uint64_t n, cudaStream_t s)
{
    if (n == 0) return 0;
    const unsigned nb = b2_div_up(n, SCAN_TILE);
    if (nb == 1) {
        scan_apply_kernel<<<1, SCAN_THREADS, 0, s>>>(data, n, nullptr);
        B2_CHECK(ctx, cudaGetLastError());
        return 0;
    }
    uint32_t *sums = nullptr;
    B2_TRY(b2_alloc(ctx, &sums, nb));
    scan_reduce_kernel<<<nb, SCAN_THREADS, 0, s>>>(data, n, sums);
    B2_CHECK(ctx, cudaGetLastError());
    int rc = b2_exclusive_scan_u32(ctx, sums, nb, s);
    if (rc == 0) {
        scan_apply_kernel<<<nb, SCAN_THREADS, 0, s>>>(data, n, sums);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; }
    }
    cudaFree(sums);
    return rc;
}
