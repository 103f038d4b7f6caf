// api.cu -- the C ABI of include/b200spmv.h: lifetime, setup (upload + device transpose),
// exec, vectors, statistics.  Multi-GPU entry points are in multi.cu.
#include <algorithm>
#include <chrono>
#include <vector>

#include "common.cuh"

static std::string g_create_error;

static uint64_t now_ns()
{
    return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(
               std::chrono::steady_clock::now().time_since_epoch()).count();
}

static void release_matrix(b200spmv_ctx *ctx)
{
    DeviceCsr &m = ctx->csr;
    b2_free(m.rowptr); b2_free(m.colind);
    { uint64_t *v = (uint64_t *)m.vals; b2_free(v); m.vals = nullptr; }
    b2_free(m.carry_row);
    b2_free(m.seg_row); b2_free(m.tile_seg); b2_free(m.warp_seg);
    { uint64_t *v = (uint64_t *)m.carry_val; b2_free(v); m.carry_val = nullptr; }
    b2_free(ctx->csc_colptr); b2_free(ctx->csc_rowind);
    { uint64_t *v = (uint64_t *)ctx->csc_vals; b2_free(v); ctx->csc_vals = nullptr; }
    b2_close_peers(ctx);
    b2_close_packed(ctx);
    { uint64_t *v = (uint64_t *)ctx->d_x; b2_free(v); ctx->d_x = nullptr; }
    { uint64_t *v = (uint64_t *)ctx->d_y; b2_free(v); ctx->d_y = nullptr; }
    { uint64_t *v = (uint64_t *)ctx->d_yin; b2_free(v); ctx->d_yin = nullptr; }
    ctx->ready = false;
}

extern "C" int b200spmv_create(b200spmv_ctx **out, int device)
{
    if (!out) return B200SPMV_EINVAL;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) +
                         " (there is no CPU fallback in this library)";
        return B200SPMV_ECUDA;
    }
    if (device < 0 || device >= ndev) { g_create_error = "device index out of range"; return B200SPMV_EINVAL; }
    b200spmv_ctx *ctx = new b200spmv_ctx();
    ctx->device = device;
    cudaDeviceProp prop;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreate(&ctx->ev0)) != cudaSuccess || (e = cudaEventCreate(&ctx->ev1)) != cudaSuccess) {
        g_create_error = cudaGetErrorString(e);
        delete ctx;
        return B200SPMV_ECUDA;
    }
    ctx->sms = prop.multiProcessorCount;
    ctx->l2_bytes = (size_t)prop.l2CacheSize;
    ctx->max_persist_bytes = (size_t)prop.persistingL2CacheMaxSize;
    // reserve the largest L2 set-aside the device offers for persisting lines: the "vector cache"
    if (ctx->max_persist_bytes) cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, ctx->max_persist_bytes);
    *out = ctx;
    return 0;
}

extern "C" void b200spmv_destroy(b200spmv_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    b2_comm_destroy(ctx);
    release_matrix(ctx);
    for (int i = 0; i < 2; i++) if (ctx->reg_ptr[i]) { cudaHostUnregister(ctx->reg_ptr[i]); cudaGetLastError(); }
    if (ctx->d_norm) cudaFree(ctx->d_norm);
    free(ctx->h_row_splits);
    cudaEventDestroy(ctx->ev0); cudaEventDestroy(ctx->ev1);
    for (int i = 0; i < Y_CHUNKS; i++) if (ctx->ev_ychunk[i]) cudaEventDestroy(ctx->ev_ychunk[i]);
    for (int i = 0; i < 64; i++) if (ctx->ev_slab[i]) cudaEventDestroy(ctx->ev_slab[i]);
    if (ctx->ev_x0) { cudaEventDestroy(ctx->ev_x0); cudaEventDestroy(ctx->ev_x1); }
    if (ctx->comm_stream) { cudaStreamDestroy(ctx->comm_stream); cudaEventDestroy(ctx->ev_xready); cudaEventDestroy(ctx->ev_xdone); cudaEventDestroy(ctx->ev_bar); }
    for (int i = 0; i < 4; i++) if (ctx->pull_stream[i]) cudaStreamDestroy(ctx->pull_stream[i]);
    if (ctx->ev_tl[0]) for (auto &e : ctx->ev_tl) cudaEventDestroy(e);
    if (ctx->d2h_stream) { cudaStreamDestroy(ctx->d2h_stream); for (int k = 0; k < Y_CHUNKS; k++) cudaEventDestroy(ctx->ev_rowchunk[k]); }
    for (int i = 0; i < 64; i++) if (ctx->ev_blk[i]) cudaEventDestroy(ctx->ev_blk[i]);
    cudaStreamDestroy(ctx->stream); cudaStreamDestroy(ctx->copy_stream);
    delete ctx;
}

extern "C" const char *b200spmv_last_error(const b200spmv_ctx *ctx)
{
    return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

extern "C" int b200spmv_set_option(b200spmv_ctx *ctx, const char *key, int64_t value)
{
    if (!ctx || !key) return B200SPMV_EINVAL;
    const std::string k(key);
    if (k == "slabs") { if (value < 0 || value > 64) return b2_fail(ctx, B200SPMV_EINVAL, "slabs must be 0..64"); ctx->opt_slabs = (int)value; }
    else if (k == "persist_bytes") ctx->opt_persist_bytes = value;
    else if (k == "persist") { if (value < -1 || value > 1) return b2_fail(ctx, B200SPMV_EINVAL, "persist must be -1 (measured), 0 or 1"); ctx->opt_persist = (int)value; }
    else if (k == "require_aligned") ctx->opt_require_aligned = (int)value;
    else if (k == "keep_csc") ctx->opt_keep_csc = (int)value;
    else if (k == "kernel") { if (value != 3 && value != 6) return b2_fail(ctx, B200SPMV_EINVAL, "kernel must be 3 or 6"); ctx->opt_kernel = (int)value; }
    else if (k == "stream_db") ctx->opt_stream_db = value != 0;
    else if (k == "stream_pf") { if (value < 0 || value > 64) return b2_fail(ctx, B200SPMV_EINVAL, "stream_pf must be 0..64"); ctx->opt_stream_pf = (int)value; }
    else if (k == "stream_occ") { if (value < 0 || value > 8) return b2_fail(ctx, B200SPMV_EINVAL, "stream_occ must be 0..8"); ctx->opt_stream_occ = (int)value; }
    else if (k == "stream_vchunks") { if (value < 1 || value > 64) return b2_fail(ctx, B200SPMV_EINVAL, "stream_vchunks must be 1..64"); ctx->opt_stream_vchunks = (int)value; }
    else if (k == "dcsr_x16") { if (value < -1 || value > (1 << 20)) return b2_fail(ctx, B200SPMV_EINVAL, "dcsr_x16 must be -1 (measured), 0 (never) or a threshold in 1/16 entries per row"); ctx->opt_dcsr_x16 = (int)value; }
    else if (k == "prefetch_tiles") ctx->opt_prefetch_tiles = (int)std::max<int64_t>(0, value);
    else if (k == "graph") ctx->opt_graph = (int)value;
    else if (k == "prefetch_x") ctx->opt_prefetch_x = (int)value;
    else if (k == "pin_host") ctx->opt_pin_host = value != 0;
    else if (k == "dcsr_maxrow") ctx->opt_dcsr_maxrow = (int)value;
    else if (k == "exchange") { if (value < 1 || value > 3) return b2_fail(ctx, B200SPMV_EINVAL, "exchange must be 1, 2 or 3"); ctx->opt_exchange = (int)value; }
    else if (k == "pull_streams") { if (value < 1 || value > 4) return b2_fail(ctx, B200SPMV_EINVAL, "pull_streams must be 1..4"); ctx->opt_pull_streams = (int)value; }
    else if (k == "own_slabs") {
        if (ctx->ready) return b2_fail(ctx, B200SPMV_ESTATE, "set own_slabs before setup");
        if (value < 0 || value > 2) return b2_fail(ctx, B200SPMV_EINVAL, "own_slabs must be 0, 1 or 2");
        ctx->opt_own_slabs = (int)value;
    }
    else if (k == "world" || k == "rank") {
        // row-block sharding without a communicator: a single SpMV needs no collective, so a
        // rank can be given its place by hand (tests: all ranks of a split, one after the other, on one GPU)
        if (ctx->ready) return b2_fail(ctx, B200SPMV_ESTATE, "set rank / world before setup");
        if (k == "world") { if (value < 1 || value > 64) return b2_fail(ctx, B200SPMV_EINVAL, "world must be 1..64"); ctx->world = (int)value; if (ctx->rank >= ctx->world) ctx->rank = 0; }
        else { if (value < 0 || value >= ctx->world) return b2_fail(ctx, B200SPMV_EINVAL, "rank must be 0..world-1"); ctx->rank = (int)value; }
    }
    else if (k == "exec_gather") ctx->opt_exec_gather = value != 0;
    else if (k == "profile") { ctx->opt_profile = (int)value; ctx->merge_kernel_ns = 0; ctx->merge_kernel_launches = 0; ctx->prof_n = 0; memset(ctx->prof_slab_ns, 0, sizeof ctx->prof_slab_ns); }
    else return b2_fail(ctx, B200SPMV_EINVAL, "unknown option");
    return 0;
}

// number of column slabs: x slice of one slab should sit in L2 next to the streams
static int choose_slabs(const b200spmv_ctx *ctx)
{
    if (ctx->opt_slabs > 0) return ctx->opt_slabs;
    const double xbytes = 8.0 * ctx->cols;
    const double l2 = (double)ctx->l2_bytes;
    if (l2 <= 0 || xbytes <= 0.6 * l2) return 1;
    // A matrix whose entries (all but 5 %) sit within L2/16 worth of columns of the diagonal gathers from a slice of x that
    // moves with the row sweep and stays in L2 whatever the size of x: slabs would only add passes over y (pure band, 16 M
    // rows: 447 GNNZ/s in one slab, 375 in three).
    if (ctx->local_nnz > 0 && ctx->band_nnz * 100 >= ctx->local_nnz * 95) return 1;
    int s = std::min(64, (int)std::ceil(xbytes / (0.45 * l2)));
    // Every slab after the first updates y once more and shortens the row pieces the kernels work on.  When an
    // L2-sized slab would hold only a few entries per row (C5: 16 entries per row over 18 slabs; C4: 32 over 9) that
    // costs more than the L2 hits save: keep about 16/3 entries per row and slab and let part of the gathers go to HBM.
    // (per NON-EMPTY row: the empty rows of a skewed block cost the kernels nothing; profiles/r01_c4_shard_sweep.txt.)
    // Since the kernels close a row with one reduction request instead of a load and a store, short row pieces are
    // cheaper than in round 1, when the rule was 8 per row and slab: banded(11)+5 at 16 M rows runs 212 / 233 / 242 /
    // 237 / 207 GNNZ/s with 1 / 2 / 3 / 4 / 6 slabs (profiles/r02_sweep_dcsr.txt) and gets 3.
    const double per_row = ctx->nonempty_rows ? (double)ctx->local_nnz / (double)ctx->nonempty_rows : 0.0;
    int by_rows = (int)(per_row * 3.0 / 16.0 + 0.5);
    // Column skew (R-MAT: 3/4 of the entries sit in the lower half of the columns): most of the uniform slabs
    // then hold far fewer entries per row than the average says, so fewer of them pay off (8 per row and slab).
    if (ctx->upper_half_nnz * 100 < ctx->local_nnz * 35) by_rows = (int)(per_row / 8.0 + 0.5), by_rows = (by_rows * 2 + 2) / 3;
    s = std::max(1, std::min(s, by_rows));
    return s;
}

// ctx->csc_* (rank-local entries, global ids) -> slab-major CSR + merge-path tiles + vectors
static int csc_to_csr(b200spmv_ctx *ctx, uint32_t index_mask)
{
    cudaStream_t s = ctx->stream;
    DeviceCsr &m = ctx->csr;
    const uint64_t t0 = now_ns();
    B2_TRY(b2_count_nonempty(ctx, ctx->local_rows, ctx->local_nnz, ctx->csc_rowind, index_mask, ctx->row_begin,
                             &ctx->nonempty_rows, s));
    {
        uint32_t mid = 0;
        B2_CHECK(ctx, cudaMemcpyAsync(&mid, ctx->csc_colptr + ctx->cols / 2, 4, cudaMemcpyDeviceToHost, s));
        B2_CHECK(ctx, cudaStreamSynchronize(s));
        ctx->upper_half_nnz = ctx->local_nnz - mid;
    }
    ctx->band_nnz = 0;
    if (ctx->opt_slabs == 0 && ctx->l2_bytes > 0 && 8.0 * ctx->cols > 0.6 * (double)ctx->l2_bytes)  // only when slabs are in question
        B2_TRY(b2_count_band(ctx, ctx->csc_colptr, ctx->csc_rowind, index_mask, ctx->cols, ctx->rows,
                             (uint32_t)std::max<uint64_t>(1, ctx->l2_bytes / 16 / 8), &ctx->band_nnz, s));
    ctx->slabs = choose_slabs(ctx);
    ctx->own_lo = ctx->own_hi = 0;
    bool own = ctx->opt_own_slabs && ctx->world > 1 && ctx->rows == ctx->cols && ctx->h_row_splits;
    if (own) {
        // slabs aligned with the ranks' blocks of x: every block is cut into the same number of slabs, so that
        // the slabs of this rank's own block can run while the other blocks are still on the wire
        const int per_block = (int)b2_div_up((uint64_t)ctx->slabs, (uint64_t)ctx->world);
        if (per_block * ctx->world > 64) own = false;
        else if (ctx->opt_own_slabs == 2) {
            // ... or only the own block is, and all the blocks left of it form one slab and all the blocks right of
            // it another: every extra slab is another pass over y (16 B per row), and on banded + random input a
            // remote block holds about one entry per row (C5 at 8 ranks: 8 passes -> 3)
            const int r = ctx->rank, W = ctx->world;
            int n = 0;
            if (r > 0) { ctx->h_slab_col[n] = 0; ctx->slab_blk_lo[n] = 0; ctx->slab_blk_hi[n] = r; n++; }
            ctx->own_lo = n;
            const uint64_t b = ctx->h_row_splits[r], w = ctx->h_row_splits[r + 1] - b;
            for (int k = 0; k < per_block; k++, n++) { ctx->h_slab_col[n] = (uint32_t)(b + w * k / per_block); ctx->slab_blk_lo[n] = r; ctx->slab_blk_hi[n] = r + 1; }
            ctx->own_hi = n;
            if (r + 1 < W) { ctx->h_slab_col[n] = ctx->h_row_splits[r + 1]; ctx->slab_blk_lo[n] = r + 1; ctx->slab_blk_hi[n] = W; n++; }
            ctx->slabs = n;
            ctx->h_slab_col[n] = ctx->cols;
            ctx->slab_cols = 0;
        } else {
            ctx->slabs = per_block * ctx->world;
            for (int r = 0; r < ctx->world; r++) {
                const uint64_t b = ctx->h_row_splits[r], w = ctx->h_row_splits[r + 1] - b;
                for (int k = 0; k < per_block; k++) {
                    ctx->h_slab_col[r * per_block + k] = (uint32_t)(b + w * k / per_block);
                    ctx->slab_blk_lo[r * per_block + k] = r;
                    ctx->slab_blk_hi[r * per_block + k] = r + 1;
                }
            }
            ctx->h_slab_col[ctx->slabs] = ctx->cols;
            ctx->own_lo = ctx->rank * per_block;
            ctx->own_hi = ctx->own_lo + per_block;
            ctx->slab_cols = 0;
        }
    }
    if (!own) {
        ctx->slab_cols = (uint32_t)(((uint64_t)b2_div_up(ctx->cols, ctx->slabs) + 31) & ~31ull);
        ctx->slabs = (int)b2_div_up(ctx->cols, ctx->slab_cols);
        for (int sl = 0; sl <= ctx->slabs; sl++)
            ctx->h_slab_col[sl] = (uint32_t)std::min<uint64_t>((uint64_t)sl * ctx->slab_cols, ctx->cols);
    }
    ctx->rowptr_stride = ((size_t)ctx->local_rows + 1 + 3) & ~(size_t)3;
    // slab entry ranges in the CSC (columns are contiguous there), each padded to a whole number
    // of flagged-kernel tiles plus at least one pad entry (the sentinel that closes the last row)
    std::vector<uint32_t> e0(ctx->slabs), e1(ctx->slabs);
    uint64_t off = 0;
    for (int sl = 0; sl < ctx->slabs; sl++) {
        const uint32_t c0 = ctx->h_slab_col[sl], c1 = ctx->h_slab_col[sl + 1];
        B2_CHECK(ctx, cudaMemcpyAsync(&e0[sl], ctx->csc_colptr + c0, 4, cudaMemcpyDeviceToHost, s));
        B2_CHECK(ctx, cudaMemcpyAsync(&e1[sl], ctx->csc_colptr + c1, 4, cudaMemcpyDeviceToHost, s));
        B2_CHECK(ctx, cudaStreamSynchronize(s));
        ctx->h_slab_off[sl] = off;
        ctx->h_slab_nnz[sl] = e1[sl] - e0[sl];
        off += ((uint64_t)(e1[sl] - e0[sl]) + B2_FLAG_TILE) / B2_FLAG_TILE * B2_FLAG_TILE;
    }
    ctx->h_slab_off[ctx->slabs] = off;
    ctx->padded_nnz = off;
    if (off >= (1ull << 32)) return b2_fail(ctx, B200SPMV_EINVAL, "padded entry count must stay below 2^32");
    B2_TRY(b2_alloc(ctx, &m.rowptr, ctx->rowptr_stride * ctx->slabs));
    B2_TRY(b2_alloc(ctx, &m.colind, off));
    { uint64_t *v = nullptr; B2_TRY(b2_alloc(ctx, &v, off)); m.vals = v; }
    // pads (between slabs and after the last entry) read as column 0 / value 0
    B2_CHECK(ctx, cudaMemsetAsync(m.colind, 0, off * 4 + 256, s));
    B2_CHECK(ctx, cudaMemsetAsync(m.vals, 0, off * 8 + 256, s));
    for (int sl = 0; sl < ctx->slabs; sl++) {
        const uint32_t o = (uint32_t)ctx->h_slab_off[sl];
        B2_TRY(b2_compress(ctx, ctx->local_rows, (uint64_t)e1[sl] - e0[sl], ctx->csc_rowind + e0[sl], index_mask,
                           ctx->row_begin, ctx->csc_colptr, ctx->cols, e0[sl], nullptr,
                           (const uint64_t *)ctx->csc_vals + e0[sl], m.rowptr + sl * ctx->rowptr_stride, o,
                           m.colind + o, (uint64_t *)m.vals + o, s));
    }
    B2_CHECK(ctx, cudaStreamSynchronize(s));
    ctx->transposetime_ns = now_ns() - t0;
    if (!ctx->opt_keep_csc) {
        b2_free(ctx->csc_colptr); b2_free(ctx->csc_rowind);
        uint64_t *v = (uint64_t *)ctx->csc_vals; b2_free(v); ctx->csc_vals = nullptr;
    }
    B2_TRY(b2_build_flags(ctx));
    { uint64_t *v = nullptr; B2_TRY(b2_alloc(ctx, &v, ctx->cols)); ctx->d_x = v; }
    { uint64_t *v = nullptr; B2_TRY(b2_alloc(ctx, &v, ctx->rows)); ctx->d_y = v; }
    B2_CHECK(ctx, cudaMemsetAsync(ctx->d_x, 0, (size_t)ctx->cols * 8, s));
    B2_CHECK(ctx, cudaMemsetAsync(ctx->d_y, 0, (size_t)ctx->rows * 8, s));
    B2_CHECK(ctx, cudaStreamSynchronize(s));
    ctx->persist_set = false;
    B2_TRY(b2_tune_slab_kernels(ctx));
    ctx->ready = true;
    return 0;
}

static bool aligned64(const void *p) { return ((uintptr_t)p & 63) == 0; }

extern "C" int b200spmv_setup_csc(b200spmv_ctx *ctx, uint32_t rows, uint32_t cols, uint32_t nnz,
                                  const uint32_t *colptr, const uint32_t *rowind, const void *vals,
                                  int dtype, uint32_t index_mask)
{
    if (!ctx) return B200SPMV_EINVAL;
    // SparseMatrix::fromMemory rejects zero dimensions (software/SparseMatrix.cpp:33-38)
    if (rows == 0 || cols == 0 || nnz == 0) return b2_fail(ctx, B200SPMV_EINVAL, "matrix data is invalid (zero rows, cols or nz)");
    if (!colptr || !rowind || !vals) return b2_fail(ctx, B200SPMV_EINVAL, "null matrix array");
    if (dtype != B200SPMV_F64 && dtype != B200SPMV_U64) return b2_fail(ctx, B200SPMV_EINVAL, "dtype must be F64 or U64");
    if (ctx->opt_require_aligned && !(aligned64(colptr) && aligned64(rowind) && aligned64(vals)))
        return b2_fail(ctx, B200SPMV_EALIGN, "matrix arrays must be 64-byte aligned (HardwareSpMV.cpp:6-16)");
    if (colptr[0] != 0 || colptr[cols] != nnz) return b2_fail(ctx, B200SPMV_EINVAL, "colptr[0] != 0 or colptr[cols] != nnz");
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    const uint64_t t0 = now_ns();
    release_matrix(ctx);
    ctx->dtype = dtype; ctx->rows = rows; ctx->cols = cols; ctx->nnz = nnz;
    cudaStream_t s = ctx->stream;

    // upload the whole CSC, then (when sharded) keep this rank's nnz-balanced row block
    uint32_t *d_colptr = nullptr, *d_rowind = nullptr; uint64_t *d_vals = nullptr;
    B2_TRY(b2_alloc(ctx, &d_colptr, (size_t)cols + 1));
    B2_TRY(b2_alloc(ctx, &d_rowind, (size_t)nnz));
    B2_TRY(b2_alloc(ctx, &d_vals, (size_t)nnz));
    B2_CHECK(ctx, cudaMemcpyAsync(d_colptr, colptr, ((size_t)cols + 1) * 4, cudaMemcpyHostToDevice, s));
    B2_CHECK(ctx, cudaMemcpyAsync(d_rowind, rowind, (size_t)nnz * 4, cudaMemcpyHostToDevice, s));
    B2_CHECK(ctx, cudaMemcpyAsync(d_vals, vals, (size_t)nnz * 8, cudaMemcpyHostToDevice, s));
    B2_CHECK(ctx, cudaStreamSynchronize(s));
    ctx->csc_colptr = d_colptr; ctx->csc_rowind = d_rowind; ctx->csc_vals = d_vals;
    {   // nothing below may index with unchecked caller data (device atomics and scatters)
        const int rc = b2_validate_indices(ctx, d_rowind, nnz, index_mask, rows, d_colptr, cols, s);
        if (rc) { release_matrix(ctx); return rc; }
    }
    ctx->row_begin = 0; ctx->row_end = rows; ctx->local_rows = rows; ctx->local_nnz = nnz;
    if (ctx->world > 1) B2_TRY(b2_shard_csc(ctx, index_mask));
    B2_TRY(csc_to_csr(ctx, index_mask));
    ctx->setuptime_ns = now_ns() - t0;
    return 0;
}

extern "C" int b200spmv_setup_synthetic(b200spmv_ctx *ctx, int kind, uint32_t rows, uint32_t cols,
                                        uint32_t nnz_per_row, uint64_t seed, int dtype)
{
    if (!ctx) return B200SPMV_EINVAL;
    if (rows == 0 || cols == 0 || nnz_per_row == 0) return b2_fail(ctx, B200SPMV_EINVAL, "matrix data is invalid (zero rows, cols or nz)");
    if (dtype != B200SPMV_F64 && dtype != B200SPMV_U64) return b2_fail(ctx, B200SPMV_EINVAL, "dtype must be F64 or U64");
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    const uint64_t t0 = now_ns();
    release_matrix(ctx);
    ctx->dtype = dtype; ctx->rows = rows; ctx->cols = cols;
    B2_TRY(b2_synth_build_csc(ctx, kind, nnz_per_row, seed));
    B2_TRY(csc_to_csr(ctx, 0xFFFFFFFFu));
    ctx->setuptime_ns = now_ns() - t0;
    return 0;
}

// ---- exec -----------------------------------------------------------------------------------

// option pin_host: page-lock a pageable host vector where it lies (once per pointer); failure is not an error,
// the copies then go through the driver's staging buffers as before
static void pin_in_place(b200spmv_ctx *ctx, int slot, const void *ptr, size_t bytes)
{
    if (!ctx->opt_pin_host || bytes < (256u << 10)) return;
    if (ctx->reg_ptr[slot] == ptr && ctx->reg_bytes[slot] >= bytes) return;
    if (ctx->reg_ptr[slot]) { cudaHostUnregister(ctx->reg_ptr[slot]); cudaGetLastError(); ctx->reg_ptr[slot] = nullptr; }
    cudaPointerAttributes pa;
    if (cudaPointerGetAttributes(&pa, ptr) == cudaSuccess && pa.type != cudaMemoryTypeUnregistered) return;  // pinned already
    cudaGetLastError();
    if (cudaHostRegister(const_cast<void *>(ptr), bytes, cudaHostRegisterDefault) == cudaSuccess) {
        ctx->reg_ptr[slot] = const_cast<void *>(ptr);
        ctx->reg_bytes[slot] = bytes;
    } else cudaGetLastError();
}

// y[i] = yin[i] (+) y[i]
template <int DT>
__global__ void vec_accumulate_kernel(uint64_t *__restrict__ y, const uint64_t *__restrict__ yin, uint64_t n)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) y[i] = Semi<DT>::add(yin[i], y[i]);
}

extern "C" int b200spmv_exec(b200spmv_ctx *ctx, const void *x_host, void *y_host, int flags)
{
    if (!ctx) return B200SPMV_EINVAL;
    if (!ctx->ready) return b2_fail(ctx, B200SPMV_ESTATE, "exec before setup");
    if (!x_host || !y_host) return b2_fail(ctx, B200SPMV_EINVAL, "null vector");
    if (ctx->opt_require_aligned && !(aligned64(x_host) && aligned64(y_host)))
        return b2_fail(ctx, B200SPMV_EALIGN, "x and y must be 64-byte aligned (HardwareSpMV.cpp:14-15)");
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const uint64_t t0 = now_ns();
    uint64_t *yh = (uint64_t *)y_host + ctx->row_begin;
    uint64_t *yd = (uint64_t *)ctx->d_y + ctx->row_begin;
    pin_in_place(ctx, 0, x_host, (size_t)ctx->cols * 8);
    pin_in_place(ctx, 1, yh, (size_t)ctx->local_rows * 8);
    // x goes up slab slice by slab slice on the copy stream: the kernels of slab s only gather from slice s,
    // so they start as soon as it has landed (b2_spmv_launch waits on ev_slab[s]) while the rest still uploads
    // Several ranks with a communicator: every rank holds the same host x, so each uploads only its 1/world of the
    // columns over its own PCIe link and the blocks are all-gathered over NVLink (ncclAllGather in place when the
    // blocks are equal).  Otherwise every rank would pull the whole vector through PCIe: at 8 ranks 8 x 134 MB of
    // host reads per step for one 134 MB vector, and the end-to-end rate FELL with the rank count (SCALE round 1).
    const bool gather_x = ctx->comm && ctx->world > 1 && ctx->opt_exec_gather && (size_t)ctx->cols * 8 >= (1u << 20);
    const bool sliced_x = !gather_x && ctx->slabs > 1 && !(flags & B200SPMV_STRICT_ORDER) && (size_t)ctx->cols * 8 >= (8u << 20);
    if (!ctx->ev_tl[0]) for (auto &e : ctx->ev_tl) B2_CHECK(ctx, cudaEventCreate(&e));
    B2_CHECK(ctx, cudaEventRecord(ctx->ev_tl[0], ctx->copy_stream));
    B2_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_tl[0], 0));
    ctx->exec_h2d_bytes = (size_t)ctx->cols * 8;
    if (gather_x) {
        uint64_t lo, hi;
        b2_exec_x_block(ctx, ctx->rank, &lo, &hi);
        if (hi > lo) B2_CHECK(ctx, cudaMemcpyAsync((uint64_t *)ctx->d_x + lo, (const uint64_t *)x_host + lo, (hi - lo) * 8, cudaMemcpyHostToDevice, s));
        B2_TRY(b2_exec_gather_x(ctx, s));
        ctx->exec_h2d_bytes = (hi - lo) * 8;
    } else if (sliced_x) {
        for (int sl = 0; sl < ctx->slabs; sl++) {
            const size_t c0 = ctx->h_slab_col[sl], c1 = ctx->h_slab_col[sl + 1];
            if (!ctx->ev_slab[sl]) B2_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_slab[sl], cudaEventDisableTiming));
            B2_CHECK(ctx, cudaMemcpyAsync((uint64_t *)ctx->d_x + c0, (const uint64_t *)x_host + c0, (c1 - c0) * 8,
                                          cudaMemcpyHostToDevice, ctx->copy_stream));
            B2_CHECK(ctx, cudaEventRecord(ctx->ev_slab[sl], ctx->copy_stream));
        }
    } else {
        B2_CHECK(ctx, cudaMemcpyAsync(ctx->d_x, x_host, (size_t)ctx->cols * 8, cudaMemcpyHostToDevice, s));
    }
    B2_CHECK(ctx, cudaEventRecord(ctx->ev_tl[1], sliced_x ? ctx->copy_stream : s));
    // Accumulating with the throughput kernels: the upload of the caller's y rides on the copy stream while
    // t = A*x is computed, then y = y_in + t (one more re-association of a sum those kernels re-associate
    // anyway).  Strict order keeps the reference's sequence ((y_in + p1) + p2) ... and uploads y first.
    const bool overlap_y = (flags & B200SPMV_ACCUMULATE) && !(flags & B200SPMV_STRICT_ORDER) && ctx->local_rows > 0 &&
                           ctx->local_rows * 8ull >= (1ull << 20);
    // Download pipeline: with kernel 3 the last slab is launched in Y_CHUNKS pieces of tiles and the rows a piece
    // completes go down (and, accumulating, have the caller's y added first) while the later pieces still run.
    // The row chunks then follow the pieces' boundaries instead of being equal.
    const bool row_chunks = ctx->local_rows * 8ull >= (1ull << 20) && b2_last_slab_chunkable(ctx, flags);
    uint64_t cb[Y_CHUNKS + 1];
    for (int k = 0; k <= Y_CHUNKS; k++) cb[k] = row_chunks ? ctx->h_chunk_row[k] : (uint64_t)ctx->local_rows * k / Y_CHUNKS;
    if (overlap_y) {
        if (!ctx->d_yin) { uint64_t *v = nullptr; B2_TRY(b2_alloc(ctx, &v, ctx->rows)); ctx->d_yin = v; }
        // in Y_CHUNKS row chunks, each with its own event: chunk k is added and sent back (other PCIe
        // direction) while the chunks after it are still on their way up
        for (int k = 0; k < Y_CHUNKS; k++) {
            const uint64_t r0 = cb[k], r1 = cb[k + 1];
            if (!ctx->ev_ychunk[k]) B2_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_ychunk[k], cudaEventDisableTiming));
            if (r1 > r0)
                B2_CHECK(ctx, cudaMemcpyAsync((uint64_t *)ctx->d_yin + ctx->row_begin + r0, yh + r0, (size_t)(r1 - r0) * 8,
                                              cudaMemcpyHostToDevice, ctx->copy_stream));
            B2_CHECK(ctx, cudaEventRecord(ctx->ev_ychunk[k], ctx->copy_stream));
        }
    } else if (flags & B200SPMV_ACCUMULATE) {
        B2_CHECK(ctx, cudaMemcpyAsync(yd, yh, (size_t)ctx->local_rows * 8, cudaMemcpyHostToDevice, s));
    }
    B2_CHECK(ctx, cudaEventRecord(ctx->ev_tl[2], overlap_y ? ctx->copy_stream : s));
    B2_CHECK(ctx, cudaEventRecord(ctx->ev0, s));
    if (row_chunks && !ctx->d2h_stream) {
        B2_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->d2h_stream, cudaStreamNonBlocking));
        for (int k = 0; k < Y_CHUNKS; k++) B2_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_rowchunk[k], cudaEventDisableTiming));
    }
    ctx->wait_slab_events = sliced_x;
    ctx->want_row_chunks = row_chunks;
    ctx->did_row_chunks = false;
    const int rc_launch = b2_spmv_launch(ctx, overlap_y ? (flags & ~B200SPMV_ACCUMULATE) : flags, s);
    ctx->wait_slab_events = false;
    ctx->want_row_chunks = false;
    B2_TRY(rc_launch);
    B2_CHECK(ctx, cudaEventRecord(ctx->ev1, s));
    const uint64_t *yin = overlap_y ? (const uint64_t *)ctx->d_yin + ctx->row_begin : nullptr;
    if (overlap_y || ctx->did_row_chunks) {
        // chunk k: (rows final) -> (caller's y landed, added) -> download; on its own stream when the kernel
        // pieces are still running on s
        cudaStream_t ds = ctx->did_row_chunks ? ctx->d2h_stream : s;
        for (int k = 0; k < Y_CHUNKS; k++) {
            const uint64_t r0 = cb[k], r1 = cb[k + 1];
            if (r1 == r0) continue;
            if (ctx->did_row_chunks) B2_CHECK(ctx, cudaStreamWaitEvent(ds, ctx->ev_rowchunk[k], 0));
            if (overlap_y) {
                B2_CHECK(ctx, cudaStreamWaitEvent(ds, ctx->ev_ychunk[k], 0));
                const unsigned grid = (unsigned)std::min<uint64_t>(b2_div_up(r1 - r0, 256), (uint64_t)ctx->sms * 16);
                if (ctx->dtype == B200SPMV_F64) vec_accumulate_kernel<B200SPMV_F64><<<grid, 256, 0, ds>>>(yd + r0, yin + r0, r1 - r0);
                else vec_accumulate_kernel<B200SPMV_U64><<<grid, 256, 0, ds>>>(yd + r0, yin + r0, r1 - r0);
                B2_CHECK(ctx, cudaGetLastError());
                ctx->launches += 1;
            }
            B2_CHECK(ctx, cudaMemcpyAsync(yh + r0, yd + r0, (size_t)(r1 - r0) * 8, cudaMemcpyDeviceToHost, ds));
        }
        B2_CHECK(ctx, cudaEventRecord(ctx->ev_tl[4], ds));
        if (ctx->did_row_chunks) B2_CHECK(ctx, cudaStreamSynchronize(ds));
    } else {
        B2_CHECK(ctx, cudaMemcpyAsync(yh, yd, (size_t)ctx->local_rows * 8, cudaMemcpyDeviceToHost, s));
        B2_CHECK(ctx, cudaEventRecord(ctx->ev_tl[4], s));
    }
    B2_CHECK(ctx, cudaStreamSynchronize(s));
    B2_CHECK(ctx, cudaStreamSynchronize(ctx->copy_stream));
    float ms = 0.f;
    for (int i = 0; i < 4; i++) {
        cudaEvent_t e = i == 2 ? ctx->ev1 : ctx->ev_tl[i < 2 ? i + 1 : 4];
        ctx->tl_ns[i] = cudaEventElapsedTime(&ms, ctx->ev_tl[0], e) == cudaSuccess ? (uint64_t)((double)ms * 1e6) : 0;
    }
    B2_CHECK(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->spmvtime_ns = (uint64_t)((double)ms * 1e6);
    const uint64_t total = now_ns() - t0;
    ctx->xfertime_ns = total > ctx->spmvtime_ns ? total - ctx->spmvtime_ns : 0;
    if (flags & B200SPMV_ACCUMULATE) ctx->exec_h2d_bytes += (uint64_t)ctx->local_rows * 8;
    ctx->exec_d2h_bytes = (uint64_t)ctx->local_rows * 8;
    return 0;
}

extern "C" int b200spmv_exec_device(b200spmv_ctx *ctx, int iters, int flags, float *ms_each)
{
    if (!ctx || iters < 1) return B200SPMV_EINVAL;
    if (!ctx->ready) return b2_fail(ctx, B200SPMV_ESTATE, "exec before setup");
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    std::vector<cudaEvent_t> ev((size_t)iters + 1);
    for (auto &e : ev) B2_CHECK(ctx, cudaEventCreate(&e));
    int rc = 0;
    // Several back-to-back SpMVs: the launches of one step (2 per slab) are captured once into a CUDA graph
    // and replayed, so a step costs one launch from the host (matters when a rank's block is small: 8 GPUs).
    cudaGraphExec_t gexec = nullptr;
    if (iters > 1 && !ctx->opt_profile && ctx->opt_graph) {
        rc = b2_spmv_prepare(ctx, flags, s);  // stream attributes are set outside the capture
        if (rc == 0) {
            cudaGraph_t g = nullptr;
            const uint64_t launches0 = ctx->launches;
            if (cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
                const int rc2 = b2_spmv_launch(ctx, flags, s);
                const cudaError_t ec = cudaStreamEndCapture(s, &g);
                if (rc2 != 0 || ec != cudaSuccess || cudaGraphInstantiate(&gexec, g, 0) != cudaSuccess) gexec = nullptr;
                if (g) cudaGraphDestroy(g);
                cudaGetLastError();
            }
            ctx->launches_per_step = ctx->launches - launches0;
            ctx->launches = launches0;  // the captured pass did not run
            cudaStreamSynchronize(s);
        }
    }
    cudaEventRecord(ev[0], s);
    for (int i = 0; i < iters && rc == 0; i++) {
        if (ctx->opt_profile && ctx->prof_n + 2 * ctx->slabs > 128) { cudaStreamSynchronize(s); b2_profile_collect(ctx); }
        if (gexec) {
            if (cudaGraphLaunch(gexec, s) != cudaSuccess) rc = b2_fail(ctx, B200SPMV_ECUDA, "graph launch failed");
            ctx->launches += ctx->launches_per_step;
        } else {
            rc = b2_spmv_launch(ctx, flags, s);
        }
        cudaEventRecord(ev[i + 1], s);
    }
    cudaError_t e = cudaStreamSynchronize(s);
    if (rc == 0 && e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; }
    if (rc == 0) {
        float last = 0.f;
        for (int i = 0; i < iters; i++) {
            cudaEventElapsedTime(&last, ev[i], ev[i + 1]);
            if (ms_each) ms_each[i] = last;
        }
        ctx->spmvtime_ns = (uint64_t)((double)last * 1e6);
    }
    for (auto &x : ev) cudaEventDestroy(x);
    if (gexec) cudaGraphExecDestroy(gexec);
    return rc;
}

#define B2_NEED_READY(ctx)                                                       \
    if (!(ctx)) return B200SPMV_EINVAL;                                          \
    if (!(ctx)->ready) return b2_fail((ctx), B200SPMV_ESTATE, "call before setup"); \
    B2_CHECK((ctx), cudaSetDevice((ctx)->device));

extern "C" int b200spmv_set_x(b200spmv_ctx *ctx, const void *x_host)
{
    B2_NEED_READY(ctx);
    B2_CHECK(ctx, cudaMemcpyAsync(ctx->d_x, x_host, (size_t)ctx->cols * 8, cudaMemcpyHostToDevice, ctx->stream));
    B2_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}
extern "C" int b200spmv_set_y(b200spmv_ctx *ctx, const void *y_host)
{
    B2_NEED_READY(ctx);
    B2_CHECK(ctx, cudaMemcpyAsync(ctx->d_y, y_host, (size_t)ctx->rows * 8, cudaMemcpyHostToDevice, ctx->stream));
    B2_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}
extern "C" int b200spmv_fill_x_synthetic(b200spmv_ctx *ctx, uint64_t seed)
{
    B2_NEED_READY(ctx);
    return b2_synth_fill_x(ctx, seed);
}
extern "C" int b200spmv_fill_y_zero(b200spmv_ctx *ctx)
{
    B2_NEED_READY(ctx);
    B2_CHECK(ctx, cudaMemsetAsync(ctx->d_y, 0, (size_t)ctx->rows * 8, ctx->stream));
    B2_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}
extern "C" int b200spmv_get_x(b200spmv_ctx *ctx, void *x_host)
{
    B2_NEED_READY(ctx);
    B2_CHECK(ctx, cudaMemcpy(x_host, ctx->d_x, (size_t)ctx->cols * 8, cudaMemcpyDeviceToHost));
    return 0;
}
extern "C" int b200spmv_get_y(b200spmv_ctx *ctx, void *y_host)
{
    B2_NEED_READY(ctx);
    B2_CHECK(ctx, cudaMemcpy((uint64_t *)y_host + ctx->row_begin, (uint64_t *)ctx->d_y + ctx->row_begin,
                             (size_t)ctx->local_rows * 8, cudaMemcpyDeviceToHost));
    return 0;
}

// CSR of this rank's rows in ORIGINAL column order (slabs re-interleaved): rowptr[local_rows+1]
// relative to the rank's first entry, colind/vals[local_nnz].
extern "C" int b200spmv_get_csr(b200spmv_ctx *ctx, uint32_t *rowptr, uint32_t *colind, void *vals)
{
    B2_NEED_READY(ctx);
    const size_t st = ctx->rowptr_stride;
    std::vector<uint32_t> rp(st * ctx->slabs), ci(ctx->padded_nnz);
    std::vector<uint64_t> v(ctx->padded_nnz);
    B2_CHECK(ctx, cudaMemcpy(rp.data(), ctx->csr.rowptr, rp.size() * 4, cudaMemcpyDeviceToHost));
    B2_CHECK(ctx, cudaMemcpy(ci.data(), ctx->csr.colind, ci.size() * 4, cudaMemcpyDeviceToHost));
    B2_CHECK(ctx, cudaMemcpy(v.data(), ctx->csr.vals, v.size() * 8, cudaMemcpyDeviceToHost));
    uint32_t p = 0;
    for (uint32_t r = 0; r < ctx->local_rows; r++) {
        rowptr[r] = p;
        for (int sl = 0; sl < ctx->slabs; sl++)
            for (uint32_t q = rp[sl * st + r]; q < rp[sl * st + r + 1]; q++) {
                colind[p] = ci[q] & ctx->colmask;
                ((uint64_t *)vals)[p] = v[q];
                p++;
            }
    }
    rowptr[ctx->local_rows] = p;
    return 0;
}

extern "C" int b200spmv_get_csc(b200spmv_ctx *ctx, uint32_t *colptr, uint32_t *rowind, void *vals)
{
    B2_NEED_READY(ctx);
    if (!ctx->csc_colptr) return b2_fail(ctx, B200SPMV_ESTATE, "CSC copy was released (set option keep_csc=1 before setup)");
    B2_CHECK(ctx, cudaMemcpy(colptr, ctx->csc_colptr, ((size_t)ctx->cols + 1) * 4, cudaMemcpyDeviceToHost));
    B2_CHECK(ctx, cudaMemcpy(rowind, ctx->csc_rowind, (size_t)ctx->local_nnz * 4, cudaMemcpyDeviceToHost));
    B2_CHECK(ctx, cudaMemcpy(vals, ctx->csc_vals, (size_t)ctx->local_nnz * 8, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int b200spmv_stat_u64(b200spmv_ctx *ctx, const char *key, uint64_t *out)
{
    if (!ctx || !key || !out) return B200SPMV_EINVAL;
    const std::string k(key);
    if (k == "rows") *out = ctx->rows;
    else if (k == "cols") *out = ctx->cols;
    else if (k == "nz") *out = ctx->nnz;
    else if (k == "local_rows") *out = ctx->local_rows;
    else if (k == "local_nz") *out = ctx->local_nnz;
    else if (k == "row_begin") *out = ctx->row_begin;
    else if (k == "row_end") *out = ctx->row_end;
    else if (k == "spmvtime_ns") *out = ctx->spmvtime_ns;
    else if (k == "xfertime_ns") *out = ctx->xfertime_ns;
    else if (k == "setuptime_ns") *out = ctx->setuptime_ns;
    else if (k == "transposetime_ns") *out = ctx->transposetime_ns;
    else if (k == "xchgtime_ns") *out = ctx->xchgtime_ns;
    else if (k == "nonempty_rows") *out = ctx->nonempty_rows;
    else if (k == "exec_x_up_ns") *out = ctx->tl_ns[0];       // timeline of the last b200spmv_exec, from its first copy
    else if (k == "exec_y_up_ns") *out = ctx->tl_ns[1];
    else if (k == "exec_kernels_ns") *out = ctx->tl_ns[2];
    else if (k == "exec_done_ns") *out = ctx->tl_ns[3];
    else if (k == "exchange_mode") *out = (uint64_t)ctx->exchange_mode;
    else if (k == "exec_h2d_bytes") *out = ctx->exec_h2d_bytes;
    else if (k == "exec_d2h_bytes") *out = ctx->exec_d2h_bytes;
    else if (k.rfind("slab_dcsr_", 0) == 0) {
        const int i = atoi(k.c_str() + 10);
        if (i < 0 || i >= ctx->slabs) return b2_fail(ctx, B200SPMV_EINVAL, "no such slab");
        *out = (uint64_t)(ctx->h_slab_dcsr[i] == 1);
    }
    else if (k == "tunetime_ns") *out = ctx->tunetime_ns;
    else if (k == "band_nz") *out = ctx->band_nnz;
    else if (k == "persist") *out = (uint64_t)(ctx->opt_persist >= 0 ? ctx->opt_persist != 0 : ctx->persist_tuned);
    else if (k.rfind("slab_ns_", 0) == 0 || k.rfind("slab_nnz_", 0) == 0 || k.rfind("slab_rows_", 0) == 0 || k.rfind("slab_maxrow_", 0) == 0) {
        const int i = atoi(k.c_str() + k.rfind('_') + 1);
        if (i < 0 || i >= ctx->slabs) return b2_fail(ctx, B200SPMV_EINVAL, "no such slab");
        *out = k[5] == 'n' && k[6] == 's' ? ctx->prof_slab_ns[i] : k[5] == 'n' ? (uint64_t)ctx->h_slab_nnz[i]
               : k[5] == 'm' ? (uint64_t)ctx->h_maxrow[i] : (uint64_t)ctx->h_nseg[i];
    }
    else if (k == "slabs") *out = (uint64_t)ctx->slabs;
    else if (k == "l2_persist_bytes") *out = ctx->max_persist_bytes;
    else if (k == "l2_bytes") *out = ctx->l2_bytes;
    else if (k == "sms") *out = (uint64_t)ctx->sms;
    else if (k == "gpus") *out = (uint64_t)ctx->world;
    else if (k == "rank") *out = (uint64_t)ctx->rank;
    else if (k == "mode") *out = (uint64_t)ctx->last_mode;
    else if (k == "launches") *out = ctx->launches;
    else if (k == "cuda_error") *out = (uint64_t)ctx->cuda_error;
    else if (k == "tiles") *out = ctx->total_tiles;
    else if (k == "merge_kernel_ns") { cudaStreamSynchronize(ctx->stream); b2_profile_collect(ctx); *out = ctx->merge_kernel_ns; }
    else if (k == "merge_kernel_launches") { cudaStreamSynchronize(ctx->stream); b2_profile_collect(ctx); *out = ctx->merge_kernel_launches; }
    else { *out = 0; return b2_fail(ctx, B200SPMV_EINVAL, "unknown stat key"); }
    return 0;
}
