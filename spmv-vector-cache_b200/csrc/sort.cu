// sort.cu -- stable LSD radix sort of (u32 key, u32 payload) pairs, 8 bits per pass.
//
// This is the device statement of the STABLE cursor scatter of software/csr2csc.c:27-33:
// sorting CSC entries by row id while keeping storage order inside a row puts each row's
// entries in exactly the order SoftwareSpMV.cpp:59-64 visits them, which strict (bit-exact
// fp64) mode needs.  A parallel atomic-cursor scatter would not be stable.
//
// Per pass: (1) per-tile digit histograms, digit-major; (2) exclusive scan over the
// histogram (scan.cu) = global base of every (digit, tile); (3) scatter with a stable
// in-tile rank: each warp owns a contiguous chunk, walks it 32 keys at a time, ranks equal
// digits with match.any, and keeps per-warp digit counters in shared memory.
#include "common.cuh"

namespace {

constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_KPT = 16;                          // keys per thread
constexpr int SORT_TILE = SORT_THREADS * SORT_KPT;    // 4096 keys per block
constexpr int RADIX = 256;

__global__ void __launch_bounds__(SORT_THREADS) sort_hist_kernel(const uint32_t *__restrict__ keys, uint64_t n,
                                                                int shift, uint32_t nblocks,
                                                                uint32_t *__restrict__ hist)
{
    __shared__ uint32_t s_cnt[RADIX];
    s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t)blockIdx.x * SORT_TILE;
#pragma unroll
    for (int j = 0; j < SORT_KPT; j++) {
        uint64_t i = base + (uint64_t)j * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&s_cnt[(keys[i] >> shift) & 0xFF], 1u);
    }
    __syncthreads();
    hist[(uint64_t)threadIdx.x * nblocks + blockIdx.x] = s_cnt[threadIdx.x];
}

template <bool IOTA>
__global__ void __launch_bounds__(SORT_THREADS)
sort_scatter_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ payload, uint64_t n,
                    int shift, uint32_t nblocks, const uint32_t *__restrict__ hist,
                    uint32_t *__restrict__ out_keys, uint32_t *__restrict__ out_payload)
{
    __shared__ uint32_t s_cnt[SORT_WARPS][RADIX];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&s_cnt[0][0])[i] = 0;
    __syncthreads();

    const uint64_t wbase = (uint64_t)blockIdx.x * SORT_TILE + (uint64_t)warp * (32 * SORT_KPT);
    uint32_t k[SORT_KPT], rank[SORT_KPT];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < SORT_KPT; j++) {
        const uint64_t i = wbase + (uint64_t)j * 32 + lane;
        const bool valid = i < n;
        k[j] = valid ? keys[i] : 0xFFFFFFFFu;
        // invalid lanes join a group of their own (digit id 256) and never touch counters
        const uint32_t d = valid ? ((k[j] >> shift) & 0xFF) : 256u;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t before = 0;
        if (valid && lane == leader) {
            before = s_cnt[warp][d];
            s_cnt[warp][d] = before + __popc(peers);
        }
        before = __shfl_sync(0xffffffffu, before, leader);
        rank[j] = before + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    {
        // thread d: exclusive scan of digit d over the warps, seeded with the global base
        const int d = threadIdx.x;
        uint32_t run = hist[(uint64_t)d * nblocks + blockIdx.x];
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) {
            uint32_t t = s_cnt[w][d];
            s_cnt[w][d] = run;
            run += t;
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < SORT_KPT; j++) {
        const uint64_t i = wbase + (uint64_t)j * 32 + lane;
        if (i < n) {
            const uint32_t d = (k[j] >> shift) & 0xFF;
            const uint32_t pos = s_cnt[warp][d] + rank[j];
            out_keys[pos] = k[j];
            out_payload[pos] = IOTA ? (uint32_t)i : payload[i];
        }
    }
}

}  // namespace

int b2_sort_pairs(b200spmv_ctx *ctx, uint32_t *keys, uint32_t *payload, uint32_t *keys_tmp,
                  uint32_t *payload_tmp, uint64_t n, int key_bits, bool payload_is_iota,
                  uint32_t **sorted_payload, cudaStream_t s)
{
    if (key_bits < 1) key_bits = 1;
    const int passes = (key_bits + 7) / 8;
    const uint32_t nblocks = b2_div_up(n, SORT_TILE);
    uint32_t *hist = nullptr;
    B2_TRY(b2_alloc(ctx, &hist, (size_t)RADIX * nblocks));
    uint32_t *src_k = keys, *src_p = payload, *dst_k = keys_tmp, *dst_p = payload_tmp;
    int rc = 0;
    for (int p = 0; p < passes && rc == 0; p++) {
        const int shift = 8 * p;
        sort_hist_kernel<<<nblocks, SORT_THREADS, 0, s>>>(src_k, n, shift, nblocks, hist);
        rc = b2_exclusive_scan_u32(ctx, hist, (uint64_t)RADIX * nblocks, s);
        if (rc) break;
        if (p == 0 && payload_is_iota)
            sort_scatter_kernel<true><<<nblocks, SORT_THREADS, 0, s>>>(src_k, src_p, n, shift, nblocks, hist, dst_k, dst_p);
        else
            sort_scatter_kernel<false><<<nblocks, SORT_THREADS, 0, s>>>(src_k, src_p, n, shift, nblocks, hist, dst_k, dst_p);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; break; }
        uint32_t *t;
        t = src_k; src_k = dst_k; dst_k = t;
        t = src_p; src_p = dst_p; dst_p = t;
    }
    cudaStreamSynchronize(s);
    cudaFree(hist);
    *sorted_payload = src_p;
    return rc;
}
