// rowmark.cu -- the row-marking ("cold miss skip") preprocessing of
// software/SparseMatrix.cpp:52-125 on the device.  The CPU versions are sequential sweeps
// with a seen-bitmap; here "first / last storage-order occurrence of each row" is a
// segmented min / max of the entry position per row id, and maxAlive is a prefix sum.
// Results are bit-identical to the reference methods (integer work).
#include <algorithm>

#include "common.cuh"

namespace {

constexpr uint32_t ROW_MASK = 0x3FFFFFFFu;  // SparseMatrix.cpp:64

__global__ void occurrence_kernel(const uint32_t *__restrict__ inds, uint64_t n, int reverse, uint32_t *__restrict__ pos)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const uint32_t r = inds[e] & ROW_MASK;
        if (reverse) atomicMax(&pos[r], (uint32_t)e); else atomicMin(&pos[r], (uint32_t)e);
    }
}
__global__ void mark_kernel(uint32_t *__restrict__ inds, uint64_t n, int shift, const uint32_t *__restrict__ pos)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const uint32_t w = inds[e];
        if (pos[w & ROW_MASK] == (uint32_t)e) inds[e] = w | (1u << shift);
    }
}
// +1 at a row's first entry, -1 (mod 2^32) at its last: SparseMatrix.cpp:101-102
__global__ void alive_delta_kernel(const uint32_t *__restrict__ inds, uint64_t n, uint32_t *__restrict__ delta)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const uint32_t w = inds[e];
        delta[e] = ((w >> 31) & 1u) - ((w >> 30) & 1u);
    }
}
// max over e of the INCLUSIVE running count (the reference takes the max after both updates)
__global__ void alive_max_kernel(const uint32_t *__restrict__ inds, const uint32_t *__restrict__ excl, uint64_t n,
                                 uint32_t *__restrict__ out)
{
    uint32_t best = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const uint32_t w = inds[e];
        best = max(best, excl[e] + ((w >> 31) & 1u) - ((w >> 30) & 1u));
    }
    for (int d = 16; d > 0; d >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, d));
    if ((threadIdx.x & 31) == 0) atomicMax(out, best);
}
// SparseMatrix.cpp:110-119 on the raw words, unsigned arithmetic; columns whose read would
// leave the array (colptr[c+1]==0 or colptr[c]==nnz) are skipped (memory safety only).
__global__ void col_span_kernel(const uint32_t *__restrict__ colptr, const uint32_t *__restrict__ inds, uint32_t cols,
                                uint32_t nnz, uint32_t *__restrict__ out)
{
    uint32_t best = 0;
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < cols) {
        const uint32_t b = colptr[c], e = colptr[c + 1];
        if (e != 0 && e <= nnz && b < nnz) best = inds[e - 1] - inds[b];
    }
    for (int d = 16; d > 0; d >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, d));
    if ((threadIdx.x & 31) == 0) atomicMax(out, best);
}

int mark_on_device(b200spmv_ctx *ctx, uint32_t rows, uint64_t n, uint32_t *d_inds, int reverse, int shift, cudaStream_t s)
{
    uint32_t *pos = nullptr;
    B2_TRY(b2_alloc(ctx, &pos, rows));
    cudaMemsetAsync(pos, reverse ? 0x00 : 0xFF, (size_t)rows * 4, s);
    const unsigned grid = (unsigned)std::min<uint64_t>(b2_div_up(n, 256), (uint64_t)ctx->sms * 32);
    occurrence_kernel<<<grid, 256, 0, s>>>(d_inds, n, reverse, pos);
    mark_kernel<<<grid, 256, 0, s>>>(d_inds, n, shift, pos);
    cudaError_t e = cudaStreamSynchronize(s);
    cudaFree(pos);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; return B200SPMV_ECUDA; }
    return 0;
}

}  // namespace

// SparseMatrix::markRowStarts(reverse, shift), SparseMatrix.cpp:52-90, in place on host inds.
// Note for reverse: a row whose LAST entry is at position 0 is still found (atomicMax from 0).
extern "C" int b200spmv_mark_row_starts(b200spmv_ctx *ctx, uint32_t rows, uint32_t nnz, uint32_t *inds,
                                        int reverse, int shift)
{
    if (!ctx || !inds || rows == 0 || nnz == 0 || shift < 0 || shift > 31) return B200SPMV_EINVAL;
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    uint32_t *d = nullptr;
    B2_TRY(b2_alloc(ctx, &d, nnz));
    int rc = 0;
    if (cudaMemcpyAsync(d, inds, (size_t)nnz * 4, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) rc = b2_fail(ctx, B200SPMV_ECUDA, "H2D failed");
    if (!rc) rc = b2_validate_indices(ctx, d, nnz, ROW_MASK, rows, nullptr, 0, ctx->stream);
    if (!rc) rc = mark_on_device(ctx, rows, nnz, d, reverse, shift, ctx->stream);
    if (!rc && (cudaMemcpyAsync(inds, d, (size_t)nnz * 4, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess)) rc = b2_fail(ctx, B200SPMV_ECUDA, "D2H failed");
    cudaFree(d);
    return rc;
}

// SparseMatrix::maxAlive, SparseMatrix.cpp:92-108: marks bit 31 / bit 30 in place (left set,
// like the reference) and returns the peak number of simultaneously "alive" rows.
extern "C" int b200spmv_max_alive(b200spmv_ctx *ctx, uint32_t rows, uint32_t nnz, uint32_t *inds, uint32_t *out)
{
    if (!ctx || !inds || !out || rows == 0 || nnz == 0) return B200SPMV_EINVAL;
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    uint32_t *d = nullptr, *delta = nullptr, *d_out = nullptr;
    int rc = 0;
    do {
        if ((rc = b2_alloc(ctx, &d, nnz))) break;
        if ((rc = b2_alloc(ctx, &delta, nnz))) break;
        if ((rc = b2_alloc(ctx, &d_out, 1))) break;
        if (cudaMemcpyAsync(d, inds, (size_t)nnz * 4, cudaMemcpyHostToDevice, s) != cudaSuccess) { rc = b2_fail(ctx, B200SPMV_ECUDA, "H2D failed"); break; }
        if ((rc = b2_validate_indices(ctx, d, nnz, ROW_MASK, rows, nullptr, 0, s))) break;
        if ((rc = mark_on_device(ctx, rows, nnz, d, 0, 31, s))) break;
        if ((rc = mark_on_device(ctx, rows, nnz, d, 1, 30, s))) break;
        const unsigned grid = (unsigned)std::min<uint64_t>(b2_div_up(nnz, 256), (uint64_t)ctx->sms * 32);
        alive_delta_kernel<<<grid, 256, 0, s>>>(d, nnz, delta);
        if ((rc = b2_exclusive_scan_u32(ctx, delta, nnz, s))) break;
        cudaMemsetAsync(d_out, 0, 4, s);
        alive_max_kernel<<<grid, 256, 0, s>>>(d, delta, nnz, d_out);
        if (cudaMemcpyAsync(out, d_out, 4, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
            cudaMemcpyAsync(inds, d, (size_t)nnz * 4, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
            cudaStreamSynchronize(s) != cudaSuccess) { rc = b2_fail(ctx, B200SPMV_ECUDA, "D2H failed"); break; }
    } while (0);
    b2_free(d); b2_free(delta); b2_free(d_out);
    return rc;
}

// SparseMatrix::maxColSpan, SparseMatrix.cpp:110-119.
extern "C" int b200spmv_max_col_span(b200spmv_ctx *ctx, uint32_t cols, uint32_t nnz, const uint32_t *colptr,
                                     const uint32_t *inds, uint32_t *out)
{
    if (!ctx || !colptr || !inds || !out || cols == 0 || nnz == 0) return B200SPMV_EINVAL;
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    uint32_t *dp = nullptr, *di = nullptr, *d_out = nullptr;
    int rc = 0;
    do {
        if ((rc = b2_alloc(ctx, &dp, (size_t)cols + 1))) break;
        if ((rc = b2_alloc(ctx, &di, nnz))) break;
        if ((rc = b2_alloc(ctx, &d_out, 1))) break;
        cudaMemcpyAsync(dp, colptr, ((size_t)cols + 1) * 4, cudaMemcpyHostToDevice, s);
        cudaMemcpyAsync(di, inds, (size_t)nnz * 4, cudaMemcpyHostToDevice, s);
        cudaMemsetAsync(d_out, 0, 4, s);
        col_span_kernel<<<b2_div_up(cols, 256), 256, 0, s>>>(dp, di, cols, nnz, d_out);
        if (cudaMemcpyAsync(out, d_out, 4, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
            cudaStreamSynchronize(s) != cudaSuccess) { rc = b2_fail(ctx, B200SPMV_ECUDA, "max_col_span failed"); break; }
    } while (0);
    b2_free(dp); b2_free(di); b2_free(d_out);
    return rc;
}
