// synth.cu -- device generators for BASELINE.json's synthetic configs (uniform-random,
// R-MAT, banded+random).  Definitions in synth_rng.h; the matrix is produced in the
// reference's CSC layout (software/SparseMatrix.h:60-69) -- rank-local entries only when the
// context is one rank of several -- and then goes through the same CSC->CSR transpose as a
// matrix loaded from a matrix directory.
#include <algorithm>

#include "common.cuh"
#include "synth_rng.h"

namespace {

__global__ void rowlen_kernel(int kind, uint32_t rows, uint32_t npr, uint32_t *__restrict__ cnt)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < rows) cnt[r] = synth_row_len(kind, r, rows, npr);
}

__global__ void rmat_hist_kernel(uint64_t seed, int scale, uint64_t nedges, uint32_t *__restrict__ cnt)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nedges; e += stride) {
        uint32_t r, c;
        synth_rmat_edge(seed, e, scale, &r, &c);
        atomicAdd(&cnt[r], 1u);
    }
}

// nnz-balanced row blocks: split k = first row whose prefix reaches k*nnz/world
__global__ void find_splits_kernel(const uint32_t *__restrict__ rowptr, uint32_t rows, int world, uint32_t *__restrict__ splits)
{
    b2_row_splits(rowptr, rows, world, splits);
}

__global__ void local_rowptr_kernel(const uint32_t *__restrict__ rowptr, uint32_t rb, uint32_t nlocal,
                                    uint32_t *__restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i <= nlocal) out[i] = rowptr[rb + i] - rowptr[rb];
}

__global__ void gen_rows_kernel(int kind, uint64_t seed, uint32_t rows, uint32_t cols, uint32_t npr, int u64,
                                uint32_t rb, uint32_t nlocal, const uint32_t *__restrict__ lrowptr,
                                uint32_t *__restrict__ col, uint64_t *__restrict__ val)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nlocal) return;
    uint32_t c[SYNTH_MAX_NPR + 2 * SYNTH_BAND_HALF + 1];
    uint64_t v[SYNTH_MAX_NPR + 2 * SYNTH_BAND_HALF + 1];
    const uint32_t n = synth_gen_row(kind, seed, rb + i, rows, cols, npr, u64, c, v);
    const uint32_t p = lrowptr[i];
    for (uint32_t k = 0; k < n; k++) { col[p + k] = c[k]; val[p + k] = v[k]; }
}

// R-MAT edges of this rank's row block, in edge order.  pos == nullptr: keep everything.
__global__ void rmat_flag_kernel(uint64_t seed, int scale, uint64_t nedges, uint32_t rb, uint32_t re,
                                 uint32_t *__restrict__ flag)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nedges; e += stride) {
        uint32_t r, c;
        synth_rmat_edge(seed, e, scale, &r, &c);
        flag[e] = (r >= rb && r < re) ? 1u : 0u;
    }
}

__global__ void rmat_emit_kernel(uint64_t seed, int scale, uint64_t nedges, int u64, uint32_t rb, uint32_t re,
                                 const uint32_t *__restrict__ pos, uint32_t *__restrict__ row,
                                 uint32_t *__restrict__ col, uint64_t *__restrict__ val)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nedges; e += stride) {
        uint32_t r, c;
        synth_rmat_edge(seed, e, scale, &r, &c);
        if (r < rb || r >= re) continue;
        const uint64_t p = pos ? pos[e] : e;
        row[p] = r; col[p] = c;
        val[p] = synth_value_bits(seed, e, 0, u64, SYNTH_RMAT, 0);
    }
}

__global__ void fill_x_kernel(uint64_t seed, uint64_t n, int u64, uint64_t *__restrict__ x)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        x[i] = synth_x_bits(seed, i, u64);
}

// ---- an independent statement of y = A*x for the synthetic matrices (u64, wrapping) ----------------------------
// Straight off the generator's definition -- no CSC, no transpose, no CSR, none of the SpMV kernels: R-MAT adds
// every edge of the rank's rows with an atomic (wrapping add is order-free), the row-generated kinds regenerate
// each row in a thread.  Used by the full-size parity tests.
__global__ void ref_rmat_kernel(uint64_t seed, int scale, uint64_t nedges, uint32_t rb, uint32_t re,
                                const uint64_t *__restrict__ x, unsigned long long *__restrict__ yref)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nedges; e += stride) {
        uint32_t r, c;
        synth_rmat_edge(seed, e, scale, &r, &c);
        if (r < rb || r >= re) continue;
        atomicAdd(&yref[r - rb], (unsigned long long)(synth_value_bits(seed, e, 0, 1, SYNTH_RMAT, 0) * x[c]));
    }
}
__global__ void ref_rows_kernel(int kind, uint64_t seed, uint32_t rows, uint32_t cols, uint32_t npr, uint32_t rb, uint32_t nlocal,
                                const uint64_t *__restrict__ x, unsigned long long *__restrict__ yref)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nlocal) return;
    uint32_t c[SYNTH_MAX_NPR + 2 * SYNTH_BAND_HALF + 1];
    uint64_t v[SYNTH_MAX_NPR + 2 * SYNTH_BAND_HALF + 1];
    const uint32_t n = synth_gen_row(kind, seed, rb + i, rows, cols, npr, 1, c, v);
    uint64_t acc = 0;
    for (uint32_t k = 0; k < n; k++) acc += v[k] * x[c[k]];
    yref[i] = acc;
}
__global__ void count_diff_kernel(const uint64_t *__restrict__ a, const uint64_t *__restrict__ b, uint64_t n,
                                  unsigned long long *__restrict__ out)
{
    unsigned c = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) c += a[i] != b[i];
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long)c);
}

}  // namespace

// mismatches = number of rows of this rank's block where the y in HBM differs from A*x recomputed from the
// generator's definition with the x in HBM (u64 only: exact)
extern "C" int b200spmv_check_synthetic_u64(b200spmv_ctx *ctx, int kind, uint32_t nnz_per_row, uint64_t seed, uint64_t *mismatches)
{
    if (!ctx || !mismatches) return B200SPMV_EINVAL;
    if (!ctx->ready) return b2_fail(ctx, B200SPMV_ESTATE, "call before setup");
    if (ctx->dtype != B200SPMV_U64) return b2_fail(ctx, B200SPMV_EINVAL, "the generator cross-check is exact arithmetic: u64 only");
    if (kind != SYNTH_RMAT && nnz_per_row > SYNTH_MAX_NPR) return b2_fail(ctx, B200SPMV_EINVAL, "nnz_per_row > 64");
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    unsigned long long *yref = nullptr, *cnt = nullptr, h = 0;
    const uint64_t nl = ctx->local_rows;
    B2_TRY(b2_alloc(ctx, &yref, nl + 1));
    cnt = yref + nl;
    B2_CHECK(ctx, cudaMemsetAsync(yref, 0, (nl + 1) * 8, s));
    if (nl) {
        if (kind == SYNTH_RMAT) {
            int scale = 0;
            while ((1ull << scale) < ctx->rows) scale++;
            ref_rmat_kernel<<<ctx->sms * 16, 256, 0, s>>>(seed, scale, (uint64_t)ctx->rows * nnz_per_row, ctx->row_begin, ctx->row_end,
                                                        (const uint64_t *)ctx->d_x, yref);
        } else {
            ref_rows_kernel<<<b2_div_up(nl, 128), 128, 0, s>>>(kind, seed, ctx->rows, ctx->cols, nnz_per_row, ctx->row_begin, (uint32_t)nl,
                                                             (const uint64_t *)ctx->d_x, yref);
        }
        count_diff_kernel<<<ctx->sms * 8, 256, 0, s>>>((const uint64_t *)yref, (const uint64_t *)ctx->d_y + ctx->row_begin, nl, cnt);
    }
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h, cnt, 8, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    b2_free(yref);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; return B200SPMV_ECUDA; }
    *mismatches = h;
    return 0;
}

int b2_synth_fill_x(b200spmv_ctx *ctx, uint64_t seed)
{
    fill_x_kernel<<<ctx->sms * 8, 256, 0, ctx->stream>>>(seed, ctx->cols, ctx->dtype == B200SPMV_U64,
                                                         (uint64_t *)ctx->d_x);
    B2_CHECK(ctx, cudaGetLastError());
    B2_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

// Fills ctx->row_begin/row_end/local_*/nnz and ctx->csc_* (rank-local CSC, global ids).
int b2_synth_build_csc(b200spmv_ctx *ctx, int kind, uint32_t npr, uint64_t seed)
{
    const uint32_t rows = ctx->rows, cols = ctx->cols;
    const int u64 = ctx->dtype == B200SPMV_U64;
    cudaStream_t s = ctx->stream;
    int scale = 0;
    uint64_t expect = (uint64_t)rows * npr;
    if (kind == SYNTH_RMAT) {
        while ((1ull << scale) < rows) scale++;
        if ((1ull << scale) != rows || rows != cols) return b2_fail(ctx, B200SPMV_EINVAL, "R-MAT needs rows = cols = 2^scale");
    } else if (kind == SYNTH_BANDED) {
        if (rows != cols || rows < 2 * SYNTH_BAND_HALF + 1) return b2_fail(ctx, B200SPMV_EINVAL, "banded matrix must be square, n >= 11");
        expect += (uint64_t)rows * (2 * SYNTH_BAND_HALF + 1) - SYNTH_BAND_HALF * (SYNTH_BAND_HALF + 1);
    } else if (kind == SYNTH_BAND) {
        if (rows != cols || rows < npr) return b2_fail(ctx, B200SPMV_EINVAL, "pure band matrix must be square, n >= nnz_per_row");
        const uint64_t h = npr / 2, u = npr - 1 - h;  // rows clipped at the top lose h, h-1, ...; at the bottom u, u-1, ...
        expect -= h * (h + 1) / 2 + u * (u + 1) / 2;
    } else if (kind != SYNTH_UNIFORM) {
        return b2_fail(ctx, B200SPMV_EINVAL, "unknown synthetic kind");
    }
    if (npr > SYNTH_MAX_NPR && kind != SYNTH_RMAT) return b2_fail(ctx, B200SPMV_EINVAL, "nnz_per_row > 64");
    if (expect >= (1ull << 32)) return b2_fail(ctx, B200SPMV_EINVAL, "nnz must stay below 2^32 (u32 numNZ, SparseMatrix.h:11)");
    ctx->nnz = expect;

    uint32_t *rowptr = nullptr, *d_splits = nullptr, *lrowptr = nullptr, *lrow = nullptr, *lcol = nullptr,
             *flag = nullptr;
    uint64_t *lval = nullptr;
    int rc = 0;
    const unsigned wide = ctx->sms * 16;
    do {
        if ((rc = b2_alloc(ctx, &rowptr, (size_t)rows + 1))) break;
        cudaMemsetAsync(rowptr, 0, ((size_t)rows + 1) * 4, s);
        if (kind == SYNTH_RMAT) rmat_hist_kernel<<<wide, 256, 0, s>>>(seed, scale, expect, rowptr);
        else rowlen_kernel<<<b2_div_up(rows, 256), 256, 0, s>>>(kind, rows, npr, rowptr);
        if ((rc = b2_exclusive_scan_u32(ctx, rowptr, (uint64_t)rows + 1, s))) break;
        uint32_t splits[65];
        if ((rc = b2_alloc(ctx, &d_splits, 65))) break;
        find_splits_kernel<<<1, 128, 0, s>>>(rowptr, rows, ctx->world, d_splits);
        if (cudaMemcpyAsync(splits, d_splits, (ctx->world + 1) * 4, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
            cudaStreamSynchronize(s) != cudaSuccess) { rc = b2_fail(ctx, B200SPMV_ECUDA, "splits copy failed"); break; }
        if (!ctx->h_row_splits) ctx->h_row_splits = (uint32_t *)malloc(65 * 4);
        memcpy(ctx->h_row_splits, splits, (ctx->world + 1) * 4);
        const uint32_t rb = splits[ctx->rank], re = splits[ctx->rank + 1];
        ctx->row_begin = rb; ctx->row_end = re; ctx->local_rows = re - rb;
        uint32_t pb = 0, pe = 0;
        cudaMemcpyAsync(&pb, rowptr + rb, 4, cudaMemcpyDeviceToHost, s);
        cudaMemcpyAsync(&pe, rowptr + re, 4, cudaMemcpyDeviceToHost, s);
        if (cudaStreamSynchronize(s) != cudaSuccess) { rc = b2_fail(ctx, B200SPMV_ECUDA, "row block copy failed"); break; }
        const uint64_t ln = pe - pb;
        ctx->local_nnz = ln;

        if ((rc = b2_alloc(ctx, &lcol, ln))) break;
        if ((rc = b2_alloc(ctx, &lval, ln))) break;
        if ((rc = b2_alloc(ctx, &ctx->csc_colptr, (size_t)cols + 1))) break;
        if ((rc = b2_alloc(ctx, &ctx->csc_rowind, ln))) break;
        { uint64_t *v = nullptr; if ((rc = b2_alloc(ctx, &v, ln))) break; ctx->csc_vals = v; }

        if (kind == SYNTH_RMAT) {
            if ((rc = b2_alloc(ctx, &lrow, ln))) break;
            if (ctx->world > 1) {
                if ((rc = b2_alloc(ctx, &flag, expect + 1))) break;
                rmat_flag_kernel<<<wide, 256, 0, s>>>(seed, scale, expect, rb, re, flag);
                if ((rc = b2_exclusive_scan_u32(ctx, flag, expect, s))) break;
            }
            rmat_emit_kernel<<<wide, 256, 0, s>>>(seed, scale, expect, u64, rb, re, flag, lrow, lcol, lval);
            b2_free(rowptr);
            // COO (edge order) -> CSC: group by column, stable
            rc = b2_compress(ctx, cols, ln, lcol, 0xFFFFFFFFu, 0, nullptr, 0, 0, lrow, lval, ctx->csc_colptr, 0,
                             ctx->csc_rowind, ctx->csc_vals, s);
        } else {
            if ((rc = b2_alloc(ctx, &lrowptr, (size_t)ctx->local_rows + 1))) break;
            local_rowptr_kernel<<<b2_div_up((uint64_t)ctx->local_rows + 1, 256), 256, 0, s>>>(rowptr, rb, ctx->local_rows, lrowptr);
            if (ctx->local_rows)
                gen_rows_kernel<<<b2_div_up(ctx->local_rows, 128), 128, 0, s>>>(kind, seed, rows, cols, npr, u64, rb,
                                                                              ctx->local_rows, lrowptr, lcol, lval);
            b2_free(rowptr);
            // CSR rows [rb, re) -> CSC with GLOBAL row ids: the other index of position p is
            // rb + (local row that owns p); b2_compress returns the local owner, so shift after.
            rc = b2_compress(ctx, cols, ln, lcol, 0xFFFFFFFFu, 0, lrowptr, ctx->local_rows, 0, nullptr, lval,
                             ctx->csc_colptr, 0, ctx->csc_rowind, ctx->csc_vals, s);
            if (rc == 0 && rb != 0) rc = b2_add_u32(ctx, ctx->csc_rowind, ln, rb, s);
        }
        cudaError_t e = cudaGetLastError();
        if (rc == 0 && e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; }
    } while (0);
    b2_free(rowptr); b2_free(d_splits); b2_free(lrowptr); b2_free(lrow); b2_free(lcol); b2_free(flag); b2_free(lval);
    return rc;
}
