// transpose.cu -- compressed-format transpose on the device (CSC -> CSR and back).
//
// Device analogue of software/csr2csc.c:11-39 with the roles of rows and columns
// selectable: count per destination segment (:20) -> running sum (:22) -> STABLE scatter
// (:27-33).  The stable scatter is done as a stable radix sort of (destination id, source
// position) pairs (sort.cu) followed by one gather, because a parallel cursor scatter
// cannot keep storage order inside a segment.
#include <algorithm>

#include "common.cuh"

namespace {

__global__ void keys_count_kernel(const uint32_t *__restrict__ key, uint32_t mask, uint32_t sub, uint64_t n,
                                  uint32_t *__restrict__ keys_out, uint32_t *__restrict__ counts)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const uint32_t k = (key[e] & mask) - sub;
        keys_out[e] = k;
        atomicAdd(&counts[k], 1u);  // csr2csc.c:20
    }
}

// segment that owns source position e: last s with ptr[s] <= e (empty segments skipped)
__device__ __forceinline__ uint32_t owner_segment(const uint32_t *__restrict__ ptr, uint32_t nseg, uint32_t e)
{
    uint32_t lo = 0, hi = nseg;  // invariant: ptr[lo] <= e < ptr[hi]
    while (hi - lo > 1) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (__ldg(ptr + mid) <= e) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void gather_kernel(const uint32_t *__restrict__ perm, uint64_t n,
                              const uint32_t *__restrict__ src_ptr, uint32_t src_major, uint32_t src_pos_off,
                              const uint32_t *__restrict__ src_other, const uint64_t *__restrict__ src_vals,
                              uint32_t *__restrict__ dst_ind, uint64_t *__restrict__ dst_vals)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride) {
        const uint32_t e = perm[p];
        dst_ind[p] = src_other ? src_other[e] : owner_segment(src_ptr, src_major, e + src_pos_off);
        if (src_vals) dst_vals[p] = src_vals[e];
    }
}

__global__ void add_u32_kernel(uint32_t *__restrict__ a, uint64_t n, uint32_t add)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) a[i] += add;
}

// out[0] = max over e of (ind[e] & mask); out[1] != 0 when ptr[0..nptr] is not non-decreasing or leaves [0, n]
__global__ void validate_kernel(const uint32_t *__restrict__ ind, uint64_t n, uint32_t mask, const uint32_t *__restrict__ ptr,
                                uint64_t nptr, uint32_t *__restrict__ out)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x, i0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t m = 0, bad = 0;
    for (uint64_t e = i0; e < n; e += stride) m = max(m, ind[e] & mask);
    if (ptr)
        for (uint64_t c = i0; c < nptr; c += stride) bad |= (ptr[c] > ptr[c + 1]) || ((uint64_t)ptr[c + 1] > n);
    m = __reduce_max_sync(0xFFFFFFFFu, m);
    bad = __reduce_or_sync(0xFFFFFFFFu, bad);
    if ((threadIdx.x & 31) == 0) {
        if (m) atomicMax(out, m);
        if (bad) atomicOr(out + 1, 1u);
    }
}

}  // namespace

// Index arrays that come from the caller are checked once before anything indexes with them: every
// (ind[e] & mask) must be below `limit` and, when given, ptr[0..nptr] must be non-decreasing within [0, n].
// A matrix that went through SparseMatrix::markRowStarts (bits 31 / 30 set, software/main.cpp:228-229) passed
// with the default mask would otherwise scatter out of bounds on the device.
int b2_validate_indices(b200spmv_ctx *ctx, const uint32_t *d_ind, uint64_t n, uint32_t mask, uint32_t limit,
                        const uint32_t *d_ptr, uint64_t nptr, cudaStream_t s)
{
    uint32_t *d_out = nullptr, h[2] = {0, 0};
    B2_TRY(b2_alloc(ctx, &d_out, 2));
    cudaError_t e = cudaMemsetAsync(d_out, 0, 8, s);
    const uint64_t work = std::max<uint64_t>(n, nptr);
    validate_kernel<<<(unsigned)std::min<uint64_t>(std::max<uint64_t>(b2_div_up(work, 256), 1), (uint64_t)ctx->sms * 32), 256, 0, s>>>(
        d_ind, n, mask, d_ptr, nptr, d_out);
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, d_out, 8, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    b2_free(d_out);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; return B200SPMV_ECUDA; }
    if (h[1]) return b2_fail(ctx, B200SPMV_EINVAL, "pointer array is not non-decreasing within [0, nnz]");
    if (h[0] >= limit) {
        char buf[256];
        snprintf(buf, sizeof buf, "index %u (after index_mask 0x%08X) is out of range [0, %u): marked indices "
                 "(SparseMatrix::markRowStarts sets bits 31/30) need index_mask 0x3FFFFFFF", h[0], mask, limit);
        return b2_fail(ctx, B200SPMV_EINVAL, buf);
    }
    return 0;
}

// rows that hold at least one of the n entries (bitmap + popcount); the slab heuristic counts entries per
// NON-EMPTY row: the empty rows of a skewed matrix cost the kernels nothing
__global__ void mark_rows_kernel(const uint32_t *__restrict__ key, uint32_t key_mask, uint32_t key_sub, uint64_t n,
                                 uint32_t *__restrict__ bitmap)
{
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t r = (key[e] & key_mask) - key_sub, bit = 1u << (r & 31);
        if (!(bitmap[r >> 5] & bit)) atomicOr(bitmap + (r >> 5), bit);  // the plain read keeps hub rows off the atomic unit
    }
}
__global__ void popcount_kernel(const uint32_t *__restrict__ bitmap, uint64_t words, unsigned long long *__restrict__ out)
{
    unsigned c = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (uint64_t)gridDim.x * blockDim.x)
        c += __popc(bitmap[i]);
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long)c);
}

int b2_count_nonempty(b200spmv_ctx *ctx, uint32_t nkeys, uint64_t n, const uint32_t *key, uint32_t key_mask,
                      uint32_t key_sub, uint64_t *out, cudaStream_t s)
{
    *out = 0;
    if (n == 0 || nkeys == 0) return 0;
    const uint64_t words = ((uint64_t)nkeys + 31) / 32;
    uint32_t *bm = nullptr;
    B2_TRY(b2_alloc(ctx, &bm, words + 2));  // the counter sits behind the bitmap, 8-byte aligned
    unsigned long long *cnt = (unsigned long long *)(bm + ((words + 1) & ~1ull));
    int rc = 0;
    cudaError_t e = cudaMemsetAsync(bm, 0, (words + 4) * 4, s);
    const unsigned grid = (unsigned)std::min<uint64_t>(b2_div_up(n, 256), (uint64_t)ctx->sms * 32);
    mark_rows_kernel<<<grid, 256, 0, s>>>(key, key_mask, key_sub, n, bm);
    popcount_kernel<<<(unsigned)std::min<uint64_t>(b2_div_up(words, 256), (uint64_t)ctx->sms * 8), 256, 0, s>>>(bm, words, cnt);
    unsigned long long h = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h, cnt, 8, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; }
    b2_free(bm);
    *out = h;
    return rc;
}

// Entries (r, c) of a CSC block whose column lies within `window` columns of the row's place on the diagonal
// (r * cols / rows): a row-ordered sweep gathers those from a slice of x that moves with it and stays in L2 whatever
// the size of x.  One warp per column (grid-stride), lanes over the column's entries.
__global__ void __launch_bounds__(256) band_count_kernel(const uint32_t *__restrict__ colptr, const uint32_t *__restrict__ rowind,
                                                         uint32_t mask, uint32_t cols, uint32_t rows, uint32_t window,
                                                         unsigned long long *__restrict__ out)
{
    const uint32_t lane = threadIdx.x & 31, nwarps = gridDim.x * (blockDim.x >> 5);
    unsigned long long mine = 0;
    for (uint32_t c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < cols; c += nwarps) {
        const uint32_t b = colptr[c], e = colptr[c + 1];
        for (uint32_t i = b + lane; i < e; i += 32) {
            const uint64_t d = (uint64_t)(rowind[i] & mask) * cols / rows;
            const uint64_t dist = d > c ? d - c : c - d;
            mine += dist <= window;
        }
    }
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_down_sync(0xffffffffu, mine, o);
    if (lane == 0 && mine) atomicAdd(out, mine);
}

int b2_count_band(b200spmv_ctx *ctx, const uint32_t *colptr, const uint32_t *rowind, uint32_t mask, uint32_t cols, uint32_t rows,
                  uint32_t window, uint64_t *out, cudaStream_t s)
{
    *out = 0;
    if (cols == 0 || rows == 0) return 0;
    unsigned long long *cnt = nullptr;
    B2_TRY(b2_alloc(ctx, &cnt, 1));
    int rc = 0;
    cudaError_t e = cudaMemsetAsync(cnt, 0, 8, s);
    band_count_kernel<<<(unsigned)ctx->sms * 8, 256, 0, s>>>(colptr, rowind, mask, cols, rows, window, cnt);
    unsigned long long h = 0;
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h, cnt, 8, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; }
    b2_free(cnt);
    *out = h;
    return rc;
}

int b2_add_u32(b200spmv_ctx *ctx, uint32_t *a, uint64_t n, uint32_t add, cudaStream_t s)
{
    if (n == 0 || add == 0) return 0;
    add_u32_kernel<<<(unsigned)std::min<uint64_t>(b2_div_up(n, 256), (uint64_t)ctx->sms * 32), 256, 0, s>>>(a, n, add);
    B2_CHECK(ctx, cudaGetLastError());
    return 0;
}

// Group n entries by key into nkeys segments, stable.
//   (key[e] & key_mask) - key_sub : destination segment of entry e (row id for CSC->CSR)
//   the entry's other index: src_other[e] (COO input) or, when src_other == nullptr, the
//   segment of the compressed source that holds position e + src_pos_off (src_ptr[src_major+1])
// Outputs dst_ptr[nkeys+1] (+ dst_ptr_add on every element), dst_ind[n], dst_vals[n] (8-byte, optional).
int b2_compress(b200spmv_ctx *ctx, uint32_t nkeys, uint64_t n, const uint32_t *key, uint32_t key_mask,
                uint32_t key_sub, const uint32_t *src_ptr, uint32_t src_major, uint32_t src_pos_off,
                const uint32_t *src_other, const void *src_vals, uint32_t *dst_ptr, uint32_t dst_ptr_add,
                uint32_t *dst_ind, void *dst_vals, cudaStream_t s)
{
    B2_CHECK(ctx, cudaMemsetAsync(dst_ptr, 0, ((size_t)nkeys + 1) * sizeof(uint32_t), s));
    if (n == 0) return b2_add_u32(ctx, dst_ptr, (uint64_t)nkeys + 1, dst_ptr_add, s);
    uint32_t *k0 = nullptr, *k1 = nullptr, *p0 = nullptr, *p1 = nullptr, *perm = nullptr;
    int rc = 0;
    do {
        if ((rc = b2_alloc(ctx, &k0, n))) break;
        if ((rc = b2_alloc(ctx, &k1, n))) break;
        if ((rc = b2_alloc(ctx, &p0, n))) break;
        if ((rc = b2_alloc(ctx, &p1, n))) break;
        const unsigned grid = (unsigned)std::min<uint64_t>(b2_div_up(n, 256), (uint64_t)ctx->sms * 32);
        keys_count_kernel<<<grid, 256, 0, s>>>(key, key_mask, key_sub, n, k0, dst_ptr);
        if ((rc = b2_exclusive_scan_u32(ctx, dst_ptr, (uint64_t)nkeys + 1, s))) break;  // csr2csc.c:22
        if ((rc = b2_add_u32(ctx, dst_ptr, (uint64_t)nkeys + 1, dst_ptr_add, s))) break;
        int key_bits = 1;
        while (key_bits < 32 && (1ull << key_bits) < (uint64_t)nkeys) key_bits++;
        if ((rc = b2_sort_pairs(ctx, k0, p0, k1, p1, n, key_bits, true, &perm, s))) break;
        gather_kernel<<<grid, 256, 0, s>>>(perm, n, src_ptr, src_major, src_pos_off, src_other,
                                           (const uint64_t *)src_vals, dst_ind, (uint64_t *)dst_vals);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; }
    } while (0);
    b2_free(k0); b2_free(k1); b2_free(p0); b2_free(p1);
    return rc;
}
