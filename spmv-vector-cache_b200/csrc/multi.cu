// multi.cu -- one context per GPU/rank: nnz-balanced row-block sharding (no reduction on y),
// and the iterated SpMV (power iteration) whose only exchange is the refresh of the
// replicated x: an all-gather of the ranks' row blocks over NVLink.
// The reference has no counterpart (single CPU + single FPGA).
//
// NCCL is bound at run time (dlopen) so that single-GPU users need no NCCL at all and a host
// process that already carries an NCCL (e.g. torch's) shares that copy.
#include <dlfcn.h>
#include <math.h>

#include <string.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace {

typedef struct { char internal[128]; } nccl_uid;
typedef int (*fn_get_uid)(nccl_uid *);
typedef int (*fn_init_rank)(ncclComm **, int, nccl_uid, int);
typedef int (*fn_destroy)(ncclComm *);
typedef int (*fn_allreduce)(const void *, void *, size_t, int, int, ncclComm *, cudaStream_t);
typedef int (*fn_bcast)(const void *, void *, size_t, int, int, ncclComm *, cudaStream_t);
typedef int (*fn_allgather)(const void *, void *, size_t, int, ncclComm *, cudaStream_t);
typedef int (*fn_group)(void);
typedef const char *(*fn_errstr)(int);

struct NcclApi {
    void *h = nullptr;
    fn_get_uid get_uid = nullptr;
    fn_init_rank init_rank = nullptr;
    fn_destroy destroy = nullptr;
    fn_allreduce allreduce = nullptr;
    fn_bcast bcast = nullptr;
    fn_allgather allgather = nullptr;
    fn_group group_start = nullptr, group_end = nullptr;
    fn_errstr errstr = nullptr;
} g_nccl;

constexpr int NCCL_UINT64 = 5, NCCL_FLOAT64 = 8, NCCL_SUM = 0;  // ncclDataType_t / ncclRedOp_t values

bool load_nccl(std::string *why)
{
    if (g_nccl.h) return true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *n : names) if ((h = dlopen(n, RTLD_NOW | RTLD_NOLOAD))) break;  // already in the process?
    if (!h) for (const char *n : names) if ((h = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!h) { *why = std::string("cannot load libnccl: ") + dlerror(); return false; }
    g_nccl.get_uid = (fn_get_uid)dlsym(h, "ncclGetUniqueId");
    g_nccl.init_rank = (fn_init_rank)dlsym(h, "ncclCommInitRank");
    g_nccl.destroy = (fn_destroy)dlsym(h, "ncclCommDestroy");
    g_nccl.allreduce = (fn_allreduce)dlsym(h, "ncclAllReduce");
    g_nccl.bcast = (fn_bcast)dlsym(h, "ncclBroadcast");
    g_nccl.allgather = (fn_allgather)dlsym(h, "ncclAllGather");
    g_nccl.group_start = (fn_group)dlsym(h, "ncclGroupStart");
    g_nccl.group_end = (fn_group)dlsym(h, "ncclGroupEnd");
    g_nccl.errstr = (fn_errstr)dlsym(h, "ncclGetErrorString");
    if (!g_nccl.get_uid || !g_nccl.init_rank || !g_nccl.destroy || !g_nccl.allreduce || !g_nccl.bcast ||
        !g_nccl.group_start || !g_nccl.group_end) { *why = "libnccl lacks a required symbol"; return false; }
    g_nccl.h = h;
    return true;
}

int nccl_fail(b200spmv_ctx *ctx, int code, const char *what)
{
    char buf[256];
    snprintf(buf, sizeof buf, "%s: %s", what, g_nccl.errstr ? g_nccl.errstr(code) : "nccl error");
    ctx->err = buf;
    return B200SPMV_ENCCL;
}
#define NCCL_TRY(ctx, call) do { int r_ = (call); if (r_ != 0) return nccl_fail(ctx, r_, #call); } while (0)

// ---- sharding of an uploaded CSC ---------------------------------------------------------
__global__ void row_hist_kernel(const uint32_t *__restrict__ rowind, uint32_t mask, uint64_t n, uint32_t *__restrict__ cnt)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride)
        atomicAdd(&cnt[rowind[e] & mask], 1u);
}
__global__ void splits_kernel(const uint32_t *__restrict__ rowptr, uint32_t rows, int world, uint32_t *__restrict__ splits)
{
    b2_row_splits(rowptr, rows, world, splits);
}
__global__ void flag_block_kernel(const uint32_t *__restrict__ rowind, uint32_t mask, uint64_t n, uint32_t rb,
                                  uint32_t re, uint32_t *__restrict__ flag)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e <= n; e += stride) {
        const uint32_t r = e < n ? (rowind[e] & mask) : 0xFFFFFFFFu;
        flag[e] = (r >= rb && r < re) ? 1u : 0u;
    }
}
__global__ void compact_kernel(const uint32_t *__restrict__ rowind, const uint64_t *__restrict__ vals, uint32_t mask,
                               uint64_t n, uint32_t rb, uint32_t re, const uint32_t *__restrict__ pos,
                               uint32_t *__restrict__ out_row, uint64_t *__restrict__ out_val)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const uint32_t raw = rowind[e], r = raw & mask;
        if (r >= rb && r < re) { out_row[pos[e]] = raw; out_val[pos[e]] = vals[e]; }
    }
}
__global__ void remap_colptr_kernel(const uint32_t *__restrict__ colptr, uint32_t cols, const uint32_t *__restrict__ pos,
                                    uint32_t *__restrict__ out)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c <= cols) out[c] = pos[colptr[c]];
}

// ---- power iteration pieces -----------------------------------------------------------------
constexpr int RED_BLOCKS = 1024, RED_THREADS = 256;
__global__ void __launch_bounds__(RED_THREADS) sumsq_partial_kernel(const double *__restrict__ y, uint64_t n,
                                                                   double *__restrict__ partial)
{
    __shared__ double s[RED_THREADS];
    double acc = 0.0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) acc += y[i] * y[i];
    s[threadIdx.x] = acc;
    __syncthreads();
    for (int w = RED_THREADS / 2; w > 0; w >>= 1) {
        if (threadIdx.x < w) s[threadIdx.x] += s[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[blockIdx.x] = s[0];
}
__global__ void __launch_bounds__(RED_THREADS) sumsq_final_kernel(const double *__restrict__ partial, int n,
                                                                 double *__restrict__ out)
{
    __shared__ double s[RED_THREADS];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += RED_THREADS) acc += partial[i];
    s[threadIdx.x] = acc;
    __syncthreads();
    for (int w = RED_THREADS / 2; w > 0; w >>= 1) {
        if (threadIdx.x < w) s[threadIdx.x] += s[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = s[0];
}
// x[block] = y[block] / sqrt(norm2)   (u64: x = y)
__global__ void scale_into_x_kernel(const uint64_t *__restrict__ y, uint64_t n, const double *__restrict__ norm2,
                                    int is_u64, uint64_t *__restrict__ x)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const double nrm = is_u64 ? 1.0 : sqrt(norm2[0]);
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        if (is_u64) x[i] = y[i];
        else x[i] = (uint64_t)__double_as_longlong(__longlong_as_double((long long)y[i]) / nrm);
    }
}
__global__ void fill_const_kernel(uint64_t *__restrict__ x, uint64_t n, uint64_t bits)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) x[i] = bits;
}

}  // namespace

extern "C" int b200spmv_comm_unique_id(void *id128)
{
    std::string why;
    if (!id128 || !load_nccl(&why)) return B200SPMV_ENCCL;
    return g_nccl.get_uid((nccl_uid *)id128) == 0 ? 0 : B200SPMV_ENCCL;
}

extern "C" int b200spmv_comm_init(b200spmv_ctx *ctx, int rank, int world, const void *id128)
{
    if (!ctx || world < 1 || world > 64 || rank < 0 || rank >= world) return B200SPMV_EINVAL;
    if (ctx->ready) return b2_fail(ctx, B200SPMV_ESTATE, "attach the communicator before setup");
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    b2_comm_destroy(ctx);
    ctx->rank = rank; ctx->world = world;
    if (world == 1) return 0;
    std::string why;
    if (!id128 || !load_nccl(&why)) return b2_fail(ctx, B200SPMV_ENCCL, why.c_str());
    nccl_uid uid;
    memcpy(&uid, id128, sizeof uid);
    NCCL_TRY(ctx, g_nccl.init_rank(&ctx->comm, world, uid, rank));
    return 0;
}

int b2_comm_destroy(b200spmv_ctx *ctx)
{
    if (ctx->comm && g_nccl.destroy) g_nccl.destroy(ctx->comm);
    ctx->comm = nullptr;
    return 0;
}

// Keep only the entries of this rank's nnz-balanced row block (stable, so the CSC order --
// hence the reference's summation order -- is preserved inside the block).
int b2_shard_csc(b200spmv_ctx *ctx, uint32_t index_mask)
{
    cudaStream_t s = ctx->stream;
    const uint32_t rows = ctx->rows, cols = ctx->cols;
    const uint64_t n = ctx->nnz;
    const unsigned wide = ctx->sms * 16;
    uint32_t *rowptr = nullptr, *d_splits = nullptr, *flag = nullptr, *new_row = nullptr, *new_colptr = nullptr;
    uint64_t *new_val = nullptr;
    int rc = 0;
    do {
        if ((rc = b2_alloc(ctx, &rowptr, (size_t)rows + 1))) break;
        cudaMemsetAsync(rowptr, 0, ((size_t)rows + 1) * 4, s);
        row_hist_kernel<<<wide, 256, 0, s>>>(ctx->csc_rowind, index_mask, n, rowptr);
        if ((rc = b2_exclusive_scan_u32(ctx, rowptr, (uint64_t)rows + 1, s))) break;
        if ((rc = b2_alloc(ctx, &d_splits, 65))) break;
        splits_kernel<<<1, 128, 0, s>>>(rowptr, rows, ctx->world, d_splits);
        if (!ctx->h_row_splits) ctx->h_row_splits = (uint32_t *)malloc(65 * 4);
        if (cudaMemcpyAsync(ctx->h_row_splits, d_splits, (ctx->world + 1) * 4, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
            cudaStreamSynchronize(s) != cudaSuccess) { rc = b2_fail(ctx, B200SPMV_ECUDA, "splits copy failed"); break; }
        const uint32_t rb = ctx->h_row_splits[ctx->rank], re = ctx->h_row_splits[ctx->rank + 1];
        if ((rc = b2_alloc(ctx, &flag, n + 1))) break;
        flag_block_kernel<<<wide, 256, 0, s>>>(ctx->csc_rowind, index_mask, n, rb, re, flag);
        if ((rc = b2_exclusive_scan_u32(ctx, flag, n + 1, s))) break;
        uint32_t ln = 0;
        // ctx->stream is non-blocking: the legacy default stream does not wait for it
        if (cudaMemcpyAsync(&ln, flag + n, 4, cudaMemcpyDeviceToHost, s) != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess) {
            rc = b2_fail(ctx, B200SPMV_ECUDA, "block size copy failed"); break;
        }
        if ((rc = b2_alloc(ctx, &new_row, ln))) break;
        if ((rc = b2_alloc(ctx, &new_val, ln))) break;
        if ((rc = b2_alloc(ctx, &new_colptr, (size_t)cols + 1))) break;
        compact_kernel<<<wide, 256, 0, s>>>(ctx->csc_rowind, (const uint64_t *)ctx->csc_vals, index_mask, n, rb, re,
                                            flag, new_row, new_val);
        remap_colptr_kernel<<<b2_div_up((uint64_t)cols + 1, 256), 256, 0, s>>>(ctx->csc_colptr, cols, flag, new_colptr);
        cudaError_t e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; break; }
        b2_free(ctx->csc_colptr); b2_free(ctx->csc_rowind);
        { uint64_t *v = (uint64_t *)ctx->csc_vals; b2_free(v); }
        ctx->csc_colptr = new_colptr; ctx->csc_rowind = new_row; ctx->csc_vals = new_val;
        new_colptr = nullptr; new_row = nullptr; new_val = nullptr;
        ctx->row_begin = rb; ctx->row_end = re; ctx->local_rows = re - rb; ctx->local_nnz = ln;
    } while (0);
    b2_free(rowptr); b2_free(d_splits); b2_free(flag); b2_free(new_row); b2_free(new_val); b2_free(new_colptr);
    return rc;
}

// Refresh the replicated x from every rank's block (in place): variable-size all-gather as
// one grouped set of broadcasts, block r rooted at rank r.
static int allgather_x(b200spmv_ctx *ctx, cudaStream_t s)
{
    if (ctx->world == 1) return 0;
    uint64_t *x = (uint64_t *)ctx->d_x;
    // equal blocks (b2_row_splits prefers them when they are balanced): one in-place all-gather
    bool equal = g_nccl.allgather != nullptr;
    const uint32_t per = ctx->h_row_splits[1] - ctx->h_row_splits[0];
    for (int r = 0; r < ctx->world && equal; r++) equal = ctx->h_row_splits[r + 1] - ctx->h_row_splits[r] == per;
    if (equal) {
        NCCL_TRY(ctx, g_nccl.allgather(x + (size_t)ctx->rank * per, x, (size_t)per, NCCL_UINT64, ctx->comm, s));
        return 0;
    }
    NCCL_TRY(ctx, g_nccl.group_start());
    for (int r = 0; r < ctx->world; r++) {
        const uint32_t b = ctx->h_row_splits[r], e = ctx->h_row_splits[r + 1];
        if (e > b) NCCL_TRY(ctx, g_nccl.bcast(x + b, x + b, (size_t)(e - b), NCCL_UINT64, r, ctx->comm, s));
    }
    NCCL_TRY(ctx, g_nccl.group_end());
    return 0;
}

// b200spmv_exec over several ranks: every rank has uploaded only columns [lo_r, hi_r) of x, lo_r = r * cols / world
// (cut to whole 32-byte units), from the caller's host vector -- the same vector on every rank -- and the other
// blocks arrive over NVLink instead of over every rank's own PCIe link.
void b2_exec_x_block(const b200spmv_ctx *ctx, int r, uint64_t *lo, uint64_t *hi)
{
    *lo = ((uint64_t)ctx->cols * (uint64_t)r / (uint64_t)ctx->world) & ~3ull;
    *hi = r + 1 == ctx->world ? ctx->cols : (((uint64_t)ctx->cols * (uint64_t)(r + 1) / (uint64_t)ctx->world) & ~3ull);
}
int b2_exec_gather_x(b200spmv_ctx *ctx, cudaStream_t s)
{
    if (ctx->world == 1 || !ctx->comm) return 0;
    uint64_t *x = (uint64_t *)ctx->d_x;
    uint64_t lo, hi;
    b2_exec_x_block(ctx, 0, &lo, &hi);
    const uint64_t per = hi - lo;
    bool equal = g_nccl.allgather != nullptr;
    for (int r = 0; r < ctx->world && equal; r++) { b2_exec_x_block(ctx, r, &lo, &hi); equal = hi - lo == per && lo == (uint64_t)r * per; }
    if (equal) {
        NCCL_TRY(ctx, g_nccl.allgather(x + (size_t)ctx->rank * per, x, (size_t)per, NCCL_UINT64, ctx->comm, s));
        return 0;
    }
    NCCL_TRY(ctx, g_nccl.group_start());
    for (int r = 0; r < ctx->world; r++) {
        b2_exec_x_block(ctx, r, &lo, &hi);
        if (hi > lo) NCCL_TRY(ctx, g_nccl.bcast(x + lo, x + lo, (size_t)(hi - lo), NCCL_UINT64, r, ctx->comm, s));
    }
    NCCL_TRY(ctx, g_nccl.group_end());
    return 0;
}

// CUDA IPC mappings of the other ranks' x (one process per GPU on one node): handles travel through the
// communicator.  Failure to map (no peer access between two GPUs) is not an error: the caller falls back to the
// NCCL all-gather and stat "exchange_mode" says so.
static int open_peers(b200spmv_ctx *ctx)
{
    if (ctx->peers_open || ctx->peers_failed) return 0;
    cudaStream_t s = ctx->stream;
    const int W = ctx->world;
    uint64_t *d_h = nullptr;
    B2_TRY(b2_alloc(ctx, &d_h, (size_t)W * 8));
    cudaIpcMemHandle_t mine;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle travels as 8 words");
    std::vector<cudaIpcMemHandle_t> all(W);
    memset(&mine, 0, sizeof mine);
    const bool got = cudaIpcGetMemHandle(&mine, ctx->d_x) == cudaSuccess;
    if (!got) { cudaGetLastError(); memset(&mine, 0, sizeof mine); }
    B2_CHECK(ctx, cudaMemcpyAsync(d_h + (size_t)ctx->rank * 8, &mine, 64, cudaMemcpyHostToDevice, s));
    NCCL_TRY(ctx, g_nccl.allgather(d_h + (size_t)ctx->rank * 8, d_h, 8, NCCL_UINT64, ctx->comm, s));
    B2_CHECK(ctx, cudaMemcpyAsync(all.data(), d_h, (size_t)W * 64, cudaMemcpyDeviceToHost, s));
    B2_CHECK(ctx, cudaStreamSynchronize(s));
    uint64_t ok = got ? 1 : 0;
    const cudaIpcMemHandle_t zero = {};
    for (int r = 0; r < W && ok; r++) {
        if (r == ctx->rank) continue;
        if (memcmp(&all[r], &zero, 64) == 0 ||
            cudaIpcOpenMemHandle(&ctx->x_peer[r], all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            cudaGetLastError();
            ctx->x_peer[r] = nullptr;
            ok = 0;
        }
    }
    // all ranks or none: a rank that could not map everything makes everybody use the all-gather
    B2_CHECK(ctx, cudaMemcpyAsync(d_h, &ok, 8, cudaMemcpyHostToDevice, s));
    NCCL_TRY(ctx, g_nccl.allreduce(d_h, d_h, 1, NCCL_UINT64, NCCL_SUM, ctx->comm, s));
    uint64_t n_ok = 0;
    B2_CHECK(ctx, cudaMemcpyAsync(&n_ok, d_h, 8, cudaMemcpyDeviceToHost, s));
    B2_CHECK(ctx, cudaStreamSynchronize(s));
    b2_free(d_h);
    if (n_ok == (uint64_t)W) ctx->peers_open = true;
    else { b2_close_peers(ctx); ctx->peers_failed = true; }
    return 0;
}

void b2_close_peers(b200spmv_ctx *ctx)
{
    for (int r = 0; r < 64; r++)
        if (ctx->x_peer[r]) { cudaIpcCloseMemHandle(ctx->x_peer[r]); ctx->x_peer[r] = nullptr; }
    ctx->peers_open = false;
    ctx->peers_failed = false;
}

extern "C" int b200spmv_power_iterate(b200spmv_ctx *ctx, int steps, int use_current_x, int flags,
                                      float *ms_total, double *last_norm)
{
    if (!ctx || steps < 1) return B200SPMV_EINVAL;
    if (!ctx->ready) return b2_fail(ctx, B200SPMV_ESTATE, "power_iterate before setup");
    if (ctx->rows != ctx->cols) return b2_fail(ctx, B200SPMV_EINVAL, "power iteration needs a square matrix");
    if (ctx->world > 1 && !ctx->h_row_splits) return b2_fail(ctx, B200SPMV_ESTATE, "no row splits");
    if (ctx->world > 1 && !ctx->comm) return b2_fail(ctx, B200SPMV_ESTATE, "iterated SpMV over several ranks needs b200spmv_comm_init");
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    const int is_u64 = ctx->dtype == B200SPMV_U64;
    const uint64_t n = ctx->rows;
    if (!ctx->d_norm) B2_CHECK(ctx, cudaMalloc((void **)&ctx->d_norm, (RED_BLOCKS + 8) * sizeof(double)));
    double *norm2 = ctx->d_norm, *partial = ctx->d_norm + 8;
    if (!use_current_x) {
        union { double d; uint64_t u; } cv;
        cv.d = 1.0 / sqrt((double)n);
        fill_const_kernel<<<ctx->sms * 8, 256, 0, s>>>((uint64_t *)ctx->d_x, n, is_u64 ? 1ull : cv.u);
    }
    const int spmv_flags = flags & ~B200SPMV_ACCUMULATE;  // y = A*x each step
    uint64_t *yb = (uint64_t *)ctx->d_y + ctx->row_begin;
    uint64_t *xb = (uint64_t *)ctx->d_x + ctx->row_begin;
    // Overlap of the exchange with the next step (setup option own_slabs).  The slabs that cover this rank's own
    // block of x only gather what scale_into_x has just written, so they run on the compute stream while the other
    // blocks are still on the wire.  exchange 1: one NCCL all-gather on a second, higher-priority stream, the other
    // slabs wait for its event.  exchange 2: the blocks are pulled one by one out of the peers' x over NVLink by
    // the copy engines (CUDA IPC mappings; a one-word all-reduce tells that every rank has written its block),
    // starting with the right-hand neighbour's, and the slabs of a block start as soon as that block has landed.
    // Strict order keeps the slabs in column order and therefore does not overlap.
    const bool overlap = ctx->world > 1 && ctx->own_hi > ctx->own_lo && !(flags & B200SPMV_STRICT_ORDER);
    const int W = ctx->world, per_block = ctx->own_hi - ctx->own_lo;
    if (overlap && !ctx->comm_stream) {
        int lo = 0, hi = 0;
        B2_CHECK(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
        B2_CHECK(ctx, cudaStreamCreateWithPriority(&ctx->comm_stream, cudaStreamNonBlocking, hi));
        for (int i = 0; i < 4; i++) B2_CHECK(ctx, cudaStreamCreateWithPriority(&ctx->pull_stream[i], cudaStreamNonBlocking, hi));
        B2_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_xready, cudaEventDisableTiming));
        B2_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_xdone, cudaEventDisableTiming));
        B2_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_bar, cudaEventDisableTiming));
        for (int r = 0; r < W; r++) B2_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_blk[r], cudaEventDisableTiming));
    }
    int mode = overlap ? ctx->opt_exchange : 0;
    if (mode == 2) { B2_TRY(open_peers(ctx)); if (!ctx->peers_open) mode = 1; }
    ctx->exchange_mode = mode;
    cudaStream_t xs = overlap ? ctx->comm_stream : s;
    if (overlap) {  // own block first, then the blocks of ranks rank+1, rank+2, ... (the order the pulls are issued in)
        int pos = 0;
        for (int k = 0; k < W; k++) {
            const int r = (ctx->rank + k) % W;
            for (int j = 0; j < per_block; j++) ctx->h_order[pos++] = r * per_block + j;
        }
    }
    uint64_t *barrier_word = (uint64_t *)(ctx->d_norm + 4);
    bool in_flight = false;
    B2_CHECK(ctx, cudaEventRecord(ctx->ev0, s));
    for (int it = 0; it < steps; it++) {
        if (overlap) {
            for (int pos = 0; pos < ctx->slabs; pos++) ctx->slab_wait[pos] = nullptr;
            if (in_flight && mode == 1) ctx->slab_wait[per_block] = ctx->ev_xdone;
            if (in_flight && mode == 2)
                for (int k = 1; k < W; k++) ctx->slab_wait[k * per_block] = ctx->ev_blk[(ctx->rank + k) % W];
            ctx->order_n = ctx->slabs;
        }
        const int rc = b2_spmv_launch(ctx, spmv_flags, s);
        ctx->order_n = 0;
        if (rc) return rc;
        if (!is_u64) {
            sumsq_partial_kernel<<<RED_BLOCKS, RED_THREADS, 0, s>>>((const double *)yb, ctx->local_rows, partial);
            sumsq_final_kernel<<<1, RED_THREADS, 0, s>>>(partial, RED_BLOCKS, norm2);
            if (ctx->world > 1) NCCL_TRY(ctx, g_nccl.allreduce(norm2, norm2, 1, NCCL_FLOAT64, NCCL_SUM, ctx->comm, s));
        }
        // peer pulls, u64 (no norm to agree on): nobody overwrites its block before everybody has pulled it
        if (mode == 2 && is_u64) NCCL_TRY(ctx, g_nccl.allreduce(barrier_word + 1, barrier_word + 1, 1, NCCL_UINT64, NCCL_SUM, ctx->comm, s));
        scale_into_x_kernel<<<ctx->sms * 8, 256, 0, s>>>(yb, ctx->local_rows, norm2, is_u64, xb);
        ctx->launches += is_u64 ? 1 : 3;
        const bool timed = it == steps - 1 && ctx->world > 1;  // exchange time of the last step -> stat "xchgtime_ns"
        if (overlap) {
            B2_CHECK(ctx, cudaEventRecord(ctx->ev_xready, s));
            B2_CHECK(ctx, cudaStreamWaitEvent(xs, ctx->ev_xready, 0));
        }
        if (timed) {
            if (!ctx->ev_x0) { B2_CHECK(ctx, cudaEventCreate(&ctx->ev_x0)); B2_CHECK(ctx, cudaEventCreate(&ctx->ev_x1)); }
            B2_CHECK(ctx, cudaEventRecord(ctx->ev_x0, xs));
        }
        if (mode == 2) {
            // every rank's block is written once this all-reduce has gone round; the next step's norm all-reduce
            // (after every rank's last slab, hence after its pulls) is what lets a rank overwrite its block again
            NCCL_TRY(ctx, g_nccl.allreduce(barrier_word, barrier_word, 1, NCCL_UINT64, NCCL_SUM, ctx->comm, xs));
            B2_CHECK(ctx, cudaEventRecord(ctx->ev_bar, xs));
            const int ns = ctx->opt_pull_streams;
            for (int i = 0; i < ns; i++) B2_CHECK(ctx, cudaStreamWaitEvent(ctx->pull_stream[i], ctx->ev_bar, 0));
            for (int k = 1; k < W; k++) {
                const int r = (ctx->rank + k) % W;
                const size_t b = ctx->h_row_splits[r], e = ctx->h_row_splits[r + 1];
                cudaStream_t ps = ctx->pull_stream[(k - 1) % ns];
                if (e > b)
                    B2_CHECK(ctx, cudaMemcpyAsync((uint64_t *)ctx->d_x + b, (const uint64_t *)ctx->x_peer[r] + b, (e - b) * 8,
                                                  cudaMemcpyDefault, ps));
                B2_CHECK(ctx, cudaEventRecord(ctx->ev_blk[r], ps));
                B2_CHECK(ctx, cudaStreamWaitEvent(xs, ctx->ev_blk[r], 0));
            }
        } else {
            B2_TRY(allgather_x(ctx, xs));
        }
        if (timed) B2_CHECK(ctx, cudaEventRecord(ctx->ev_x1, xs));
        if (overlap) {
            B2_CHECK(ctx, cudaEventRecord(ctx->ev_xdone, xs));
            in_flight = true;
        }
    }
    if (in_flight) B2_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_xdone, 0));  // x is whole again when the call returns
    B2_CHECK(ctx, cudaEventRecord(ctx->ev1, s));
    B2_CHECK(ctx, cudaStreamSynchronize(s));
    float ms = 0.f;
    B2_CHECK(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    if (ms_total) *ms_total = ms;
    if (ctx->world > 1 && ctx->ev_x0) {
        float mx = 0.f;
        if (cudaEventElapsedTime(&mx, ctx->ev_x0, ctx->ev_x1) == cudaSuccess) ctx->xchgtime_ns = (uint64_t)((double)mx * 1e6);
    }
    if (last_norm) {
        double h = 0.0;
        if (!is_u64) B2_CHECK(ctx, cudaMemcpy(&h, norm2, 8, cudaMemcpyDeviceToHost));
        *last_norm = is_u64 ? 0.0 : sqrt(h);
    }
    return 0;
}
