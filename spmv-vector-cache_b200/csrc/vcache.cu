// vcache.cu -- the reference's namesake as a measurable GPU experiment: what would the
// FPGA's direct-mapped write-back "vector cache" over y do on this matrix?
//
// Functional model of chisel/cache-new/NoWMVectorCache.scala: ocmDepth lines of one vector element,
// index = low log2(depth) bits of the row id, tag = the rest (:14-24); all lines invalid after the fill
// phase (:197-206); a read with a matching valid tag hits (:94), anything else is a read miss (:95) that
// replaces the line, writing the old element back to DRAM first when the line was valid (the conflict
// eviction, :264-268); with cold-miss skip the first touch of a row (bit 31 of the index, set by
// SparseMatrix::markRowStarts) does not read DRAM (:240-252); the flush phase writes every valid line
// back (:208-231).  The counters are the ones software/HardwareSpMVNewCache.cpp:189-204 reports.
// Parity unpinned: the reference holds no counter fixtures and its simulator cannot run here; the tests compare this
// parallel form with a sequential restatement of the same reading of the RTL (oracle_vector_cache_model), not with the RTL.
//
// A direct-mapped cache is independent per set, so the sequential simulation parallelises: stable radix
// sort of the row-index stream by set (sort.cu), then an access misses iff it is the first of its set or
// its predecessor in the set carried a different row id.
#include <algorithm>

#include "common.cuh"

namespace {

__global__ void vc_keys_kernel(const uint32_t *__restrict__ inds, uint64_t n, uint32_t set_mask, uint32_t *__restrict__ key)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) key[e] = inds[e] & set_mask;
}

// acc[0] misses, [1] conflict evictions, [2] cold-miss skips, [3] lines valid at the end
__global__ void __launch_bounds__(256) vc_count_kernel(const uint32_t *__restrict__ sorted, uint64_t n, uint32_t set_mask,
                                                       uint32_t addr_mask, int cms, unsigned long long *__restrict__ acc)
{
    unsigned long long miss = 0, conflict = 0, cold = 0, lines = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride) {
        const uint32_t raw = sorted[p], a = raw & addr_mask;
        const bool first = p == 0 || ((sorted[p - 1] & set_mask) != (a & set_mask));
        const bool m = first || ((sorted[p - 1] & addr_mask) != a);
        miss += m;
        conflict += m && !first;
        cold += m && cms && (raw >> 31);
        lines += first;
    }
    for (int d = 16; d > 0; d >>= 1) {
        miss += __shfl_xor_sync(0xffffffffu, miss, d);
        conflict += __shfl_xor_sync(0xffffffffu, conflict, d);
        cold += __shfl_xor_sync(0xffffffffu, cold, d);
        lines += __shfl_xor_sync(0xffffffffu, lines, d);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&acc[0], miss); atomicAdd(&acc[1], conflict); atomicAdd(&acc[2], cold); atomicAdd(&acc[3], lines);
    }
}

}  // namespace

// inds: the CSC row-index stream in storage order (host), raw: with cms != 0 bit 31 is the row-start mark
// and is not part of the address (addrBits = ptrWidth-1, NoWMVectorCache.scala:15).  depth_log2 =
// log2(ocmDepth).  out[7]: reads, readMisses, conflictEvictions, coldMissSkips, ddrReads, flushWrites, hits.
extern "C" int b200spmv_vector_cache_model(b200spmv_ctx *ctx, uint32_t nnz, const uint32_t *inds, int depth_log2,
                                           int cms, uint64_t *out)
{
    if (!ctx || !inds || !out || nnz == 0 || depth_log2 < 0 || depth_log2 > 30) return B200SPMV_EINVAL;
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const uint32_t set_mask = (1u << depth_log2) - 1u, addr_mask = cms ? 0x7FFFFFFFu : 0xFFFFFFFFu;
    uint32_t *d_in = nullptr, *key = nullptr, *key2 = nullptr, *pay2 = nullptr, *sorted = nullptr;
    unsigned long long *acc = nullptr;
    int rc = 0;
    do {
        if ((rc = b2_alloc(ctx, &d_in, nnz)) || (rc = b2_alloc(ctx, &key, nnz)) || (rc = b2_alloc(ctx, &key2, nnz)) ||
            (rc = b2_alloc(ctx, &pay2, nnz)) || (rc = b2_alloc(ctx, &acc, 4))) break;
        if (cudaMemcpyAsync(d_in, inds, (size_t)nnz * 4, cudaMemcpyHostToDevice, s) != cudaSuccess) { rc = b2_fail(ctx, B200SPMV_ECUDA, "H2D failed"); break; }
        cudaMemsetAsync(acc, 0, 32, s);
        const unsigned grid = (unsigned)std::min<uint64_t>(b2_div_up(nnz, 256), (uint64_t)ctx->sms * 32);
        vc_keys_kernel<<<grid, 256, 0, s>>>(d_in, nnz, set_mask, key);
        sorted = d_in;
        if (depth_log2 > 0 && (rc = b2_sort_pairs(ctx, key, d_in, key2, pay2, nnz, depth_log2, false, &sorted, s))) break;
        vc_count_kernel<<<grid, 256, 0, s>>>(sorted, nnz, set_mask, addr_mask, cms, acc);
        unsigned long long h[4];
        if (cudaMemcpyAsync(h, acc, 32, cudaMemcpyDeviceToHost, s) != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess) {
            rc = b2_fail(ctx, B200SPMV_ECUDA, "vector cache model failed"); break;
        }
        out[0] = nnz; out[1] = h[0]; out[2] = h[1]; out[3] = h[2]; out[4] = h[0] - h[2]; out[5] = h[3]; out[6] = nnz - h[0];
    } while (0);
    b2_free(d_in); b2_free(key); b2_free(key2); b2_free(pay2); b2_free(acc);
    return rc;
}
