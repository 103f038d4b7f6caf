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
// ---- packed exchange (exchange 3) ----------------------------------------------------------------
// which rank's block column c lies in (splits[0..world], ascending)
__device__ __forceinline__ int block_of(const uint32_t *__restrict__ splits, int world, uint32_t c)
{
    int lo = 0, hi = world - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(splits + mid) <= c) lo = mid; else hi = mid - 1;
    }
    return lo;
}
// bit (c - first column of c's block) of that block's plane = some entry of this slab gathers column c
__global__ void need_mark_kernel(const uint32_t *__restrict__ colind, uint64_t n, const uint32_t *__restrict__ splits, int world,
                                 uint32_t words, uint32_t *__restrict__ bits)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint32_t g = colind[i] & 0x7FFFFFFFu;
        const int r = block_of(splits, world, g);
        const uint32_t c = g - __ldg(splits + r);
        atomicOr(&bits[(size_t)r * words + (c >> 5)], 1u << (c & 31));
    }
}
__global__ void popc_kernel(const uint32_t *__restrict__ bits, uint64_t n, uint32_t *__restrict__ out)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __popc(bits[i]);
}
// column -> position in d_unpack: start of its block's packed copy + rank among the block's needed columns
// (row-start flag kept); pads -> position 0 of the slab's first block
__global__ void remap_packed_kernel(uint32_t *__restrict__ colind, uint64_t nreal, uint64_t ntotal, const uint32_t *__restrict__ splits,
                                    int world, uint32_t words, const uint32_t *__restrict__ bits, const uint32_t *__restrict__ pref,
                                    const uint64_t *__restrict__ unpack_off, uint32_t pad_pos)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < ntotal; i += stride) {
        const uint32_t v = colind[i];
        uint32_t pos = pad_pos;
        if (i < nreal) {
            const uint32_t g = v & 0x7FFFFFFFu;
            const int r = block_of(splits, world, g);
            const uint32_t c = g - __ldg(splits + r), w = c >> 5;
            const size_t k = (size_t)r * words + w;
            pos = (uint32_t)unpack_off[r] + pref[k] + __popc(bits[k] & ((1u << (c & 31)) - 1u));
        }
        colind[i] = (v & 0x80000000u) | pos;
    }
}
// One warp per 32 columns of the rank's own block: lane = column.  For every peer whose bitmap has the lane's bit
// the value goes to that peer's packed block, at (prefix of the word) + (set bits below the lane).
__global__ void __launch_bounds__(256) pack_kernel(const uint64_t *__restrict__ xb, uint32_t width, int world, int rank,
                                                   const uint32_t *__restrict__ bits, const uint32_t *__restrict__ pref,
                                                   uint32_t words, uint64_t *__restrict__ pack, const uint64_t *__restrict__ pack_off)
{
    const uint32_t lane = threadIdx.x & 31, below = (1u << lane) - 1u;
    const uint32_t nwarps = gridDim.x * (blockDim.x >> 5);
    for (uint32_t w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < words; w += nwarps) {
        const uint32_t c = w * 32 + lane;
        const uint64_t xv = c < width ? xb[c] : 0ull;
        for (int r = 0; r < world; r++) {
            if (r == rank) continue;
            const uint32_t b = __ldg(bits + (size_t)r * words + w);
            if (b >> lane & 1u) pack[pack_off[r] + __ldg(pref + (size_t)r * words + w) + __popc(b & below)] = xv;
        }
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

void b2_close_packed(b200spmv_ctx *ctx)
{
    for (int r = 0; r < 64; r++)
        if (ctx->pack_peer[r]) { cudaIpcCloseMemHandle(ctx->pack_peer[r]); ctx->pack_peer[r] = nullptr; }
    b2_free(ctx->colind_packed); b2_free(ctx->d_unpack); b2_free(ctx->d_pack); b2_free(ctx->d_pack_off);
    b2_free(ctx->d_unpack_off); b2_free(ctx->d_splits);
    b2_free(ctx->d_pack_bits); b2_free(ctx->d_pack_pref);
    ctx->packed_ready = ctx->packed_failed = ctx->packed_active = false;
}

// Packed exchange, built once per matrix (own_slabs layout, peers already mapped):
//  1. mark, per remote block, the columns this rank's entries gather (bitmap over the block);
//  2. all-gather the bitmaps: the owner of a block learns what every peer needs of it;
//  3. prefix sums of the popcounts on both sides give a column's position in the packed block;
//  4. a second copy of colind with the remote slabs renumbered to those positions;
//  5. send buffers (mapped by the peers through CUDA IPC) and receive buffers.
// Any failure to map leaves packed_failed set and the caller falls back to whole-block pulls.
static int open_packed(b200spmv_ctx *ctx)
{
    if (ctx->packed_ready || ctx->packed_failed) return 0;
    cudaStream_t s = ctx->stream;
    const int W = ctx->world;
    uint32_t words = 0;
    for (int r = 0; r < W; r++) words = std::max(words, b2_div_up(ctx->h_row_splits[r + 1] - ctx->h_row_splits[r], 32));
    words = (words + 1) & ~1u;  // travels as 64-bit words
    ctx->pack_words = words;
    const size_t plane = (size_t)W * words;
    uint32_t *need_bits = nullptr, *need_pref = nullptr, *all_bits = nullptr;
    uint64_t *d_tab = nullptr;
    int rc = 0;
#define PK_TRY(x) do { rc = (x); if (rc) goto done; } while (0)
#define PK_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { ctx->err = cudaGetErrorString(e_); ctx->cuda_error = (int)e_; rc = B200SPMV_ECUDA; goto done; } } while (0)
    {
        PK_TRY(b2_alloc(ctx, &need_bits, plane));
        PK_TRY(b2_alloc(ctx, &need_pref, plane));
        PK_TRY(b2_alloc(ctx, &all_bits, plane * W));
        PK_TRY(b2_alloc(ctx, &ctx->d_pack_bits, plane));
        PK_TRY(b2_alloc(ctx, &ctx->d_pack_pref, plane));
        PK_TRY(b2_alloc(ctx, &d_tab, (size_t)W * W + 8 * W));
        PK_TRY(b2_alloc(ctx, &ctx->d_splits, (size_t)W + 1));
        PK_CUDA(cudaMemcpyAsync(ctx->d_splits, ctx->h_row_splits, ((size_t)W + 1) * 4, cudaMemcpyHostToDevice, s));
        PK_CUDA(cudaMemsetAsync(need_bits, 0, plane * 4, s));
        for (int sl = 0; sl < ctx->slabs; sl++) {
            if ((sl >= ctx->own_lo && sl < ctx->own_hi) || ctx->h_slab_nnz[sl] == 0) continue;
            need_mark_kernel<<<ctx->sms * 8, 256, 0, s>>>(ctx->csr.colind + ctx->h_slab_off[sl], ctx->h_slab_nnz[sl], ctx->d_splits, W,
                                                          words, need_bits);
        }
        PK_CUDA(cudaGetLastError());
        // rank q's bitmap of block r lands at all_bits[q][r]; mine to pack from are all_bits[q][rank]
        PK_CUDA(cudaMemcpyAsync(all_bits + (size_t)ctx->rank * plane, need_bits, plane * 4, cudaMemcpyDeviceToDevice, s));
        if (g_nccl.allgather(all_bits + (size_t)ctx->rank * plane, all_bits, plane / 2, NCCL_UINT64, ctx->comm, s) != 0) { rc = b2_fail(ctx, B200SPMV_ENCCL, "ncclAllGather of the column bitmaps failed"); goto done; }
        for (int q = 0; q < W; q++)
            PK_CUDA(cudaMemcpyAsync(ctx->d_pack_bits + (size_t)q * words, all_bits + (size_t)q * plane + (size_t)ctx->rank * words,
                                    (size_t)words * 4, cudaMemcpyDeviceToDevice, s));
        PK_CUDA(cudaMemsetAsync(ctx->d_pack_bits + (size_t)ctx->rank * words, 0, (size_t)words * 4, s));
        // prefix sums, block by block, on both sides; totals = last prefix + last popcount
        std::vector<uint32_t> last(4 * W);
        for (int side = 0; side < 2; side++) {
            const uint32_t *bits = side ? ctx->d_pack_bits : need_bits;
            uint32_t *pref = side ? ctx->d_pack_pref : need_pref;
            popc_kernel<<<b2_div_up(plane, 256), 256, 0, s>>>(bits, plane, pref);
            PK_CUDA(cudaGetLastError());
            for (int r = 0; r < W; r++) {
                PK_CUDA(cudaMemcpyAsync(&last[side * 2 * W + r], pref + (size_t)r * words + words - 1, 4, cudaMemcpyDeviceToHost, s));
                PK_TRY(b2_exclusive_scan_u32(ctx, pref + (size_t)r * words, words, s));
                PK_CUDA(cudaMemcpyAsync(&last[side * 2 * W + W + r], pref + (size_t)r * words + words - 1, 4, cudaMemcpyDeviceToHost, s));
            }
        }
        PK_CUDA(cudaStreamSynchronize(s));
        uint64_t uoff = 0, poff = 0;
        for (int r = 0; r < W; r++) {
            ctx->need_cnt[r] = (uint64_t)last[r] + last[W + r];
            ctx->pack_cnt[r] = (uint64_t)last[2 * W + r] + last[3 * W + r];
            ctx->unpack_off[r] = uoff;
            ctx->pack_off[r] = poff;
            uoff += (ctx->need_cnt[r] + 32 + 31) & ~31ull;  // 256-byte aligned, never empty (pads gather position 0)
            poff += (ctx->pack_cnt[r] + 32 + 31) & ~31ull;
        }
        ctx->unpack_off[W] = uoff;
        ctx->pack_off[W] = poff;
        PK_TRY(b2_alloc(ctx, &ctx->d_unpack, uoff));
        PK_TRY(b2_alloc(ctx, &ctx->d_pack, poff));
        if (uoff >= (1ull << 31)) { rc = b2_fail(ctx, B200SPMV_EINVAL, "packed blocks must stay below 2^31 values"); goto done; }
        PK_TRY(b2_alloc(ctx, &ctx->d_pack_off, (size_t)W + 1));
        PK_TRY(b2_alloc(ctx, &ctx->d_unpack_off, (size_t)W + 1));
        PK_CUDA(cudaMemsetAsync(ctx->d_unpack, 0, uoff * 8, s));
        PK_CUDA(cudaMemsetAsync(ctx->d_pack, 0, poff * 8, s));
        PK_CUDA(cudaMemcpyAsync(ctx->d_pack_off, ctx->pack_off, ((size_t)W + 1) * 8, cudaMemcpyHostToDevice, s));
        PK_CUDA(cudaMemcpyAsync(ctx->d_unpack_off, ctx->unpack_off, ((size_t)W + 1) * 8, cudaMemcpyHostToDevice, s));
        // the renumbered copy of colind
        PK_TRY(b2_alloc(ctx, &ctx->colind_packed, (size_t)ctx->padded_nnz));
        PK_CUDA(cudaMemcpyAsync(ctx->colind_packed, ctx->csr.colind, (size_t)ctx->padded_nnz * 4 + 256, cudaMemcpyDeviceToDevice, s));
        for (int sl = 0; sl < ctx->slabs; sl++) {
            if (sl >= ctx->own_lo && sl < ctx->own_hi) continue;
            remap_packed_kernel<<<ctx->sms * 8, 256, 0, s>>>(ctx->colind_packed + ctx->h_slab_off[sl], ctx->h_slab_nnz[sl],
                                                             ctx->h_slab_off[sl + 1] - ctx->h_slab_off[sl], ctx->d_splits, W, words,
                                                             need_bits, need_pref, ctx->d_unpack_off,
                                                             (uint32_t)ctx->unpack_off[ctx->slab_blk_lo[sl]]);
        }
        PK_CUDA(cudaGetLastError());
        // every rank's table of offsets and counts, and the IPC handle of its send buffer
        uint64_t *mine = d_tab + (size_t)ctx->rank * (W + 8);
        std::vector<uint64_t> row(W + 8, 0), tab((size_t)W * (W + 8));
        for (int r = 0; r < W; r++) row[r] = ctx->pack_off[r] << 32 | ctx->pack_cnt[r];
        cudaIpcMemHandle_t h;
        memset(&h, 0, sizeof h);
        const bool got = cudaIpcGetMemHandle(&h, ctx->d_pack) == cudaSuccess;
        if (!got) { cudaGetLastError(); memset(&h, 0, sizeof h); }
        memcpy(&row[W], &h, 64);
        PK_CUDA(cudaMemcpyAsync(mine, row.data(), (size_t)(W + 8) * 8, cudaMemcpyHostToDevice, s));
        if (g_nccl.allgather(mine, d_tab, (size_t)W + 8, NCCL_UINT64, ctx->comm, s) != 0) { rc = b2_fail(ctx, B200SPMV_ENCCL, "ncclAllGather of the pack tables failed"); goto done; }
        PK_CUDA(cudaMemcpyAsync(tab.data(), d_tab, tab.size() * 8, cudaMemcpyDeviceToHost, s));
        PK_CUDA(cudaStreamSynchronize(s));
        uint64_t ok = got ? 1 : 0;
        const cudaIpcMemHandle_t zero = {};
        for (int r = 0; r < W && ok; r++) {
            if (r == ctx->rank) continue;
            const uint64_t *t = &tab[(size_t)r * (W + 8)];
            ctx->peer_pack_off[r] = t[ctx->rank] >> 32;
            if ((t[ctx->rank] & 0xFFFFFFFFull) != ctx->need_cnt[r]) { ok = 0; break; }  // both sides must count the same columns
            cudaIpcMemHandle_t ph;
            memcpy(&ph, &t[W], 64);
            if (memcmp(&ph, &zero, 64) == 0 || cudaIpcOpenMemHandle(&ctx->pack_peer[r], ph, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                ctx->pack_peer[r] = nullptr;
                ok = 0;
            }
        }
        PK_CUDA(cudaMemcpyAsync(d_tab, &ok, 8, cudaMemcpyHostToDevice, s));
        if (g_nccl.allreduce(d_tab, d_tab, 1, NCCL_UINT64, NCCL_SUM, ctx->comm, s) != 0) { rc = b2_fail(ctx, B200SPMV_ENCCL, "ncclAllReduce failed"); goto done; }
        uint64_t n_ok = 0;
        PK_CUDA(cudaMemcpyAsync(&n_ok, d_tab, 8, cudaMemcpyDeviceToHost, s));
        PK_CUDA(cudaStreamSynchronize(s));
        if (n_ok == (uint64_t)W) ctx->packed_ready = true;
    }
done:
#undef PK_TRY
#undef PK_CUDA
    b2_free(need_bits); b2_free(need_pref); b2_free(all_bits); b2_free(d_tab);
    if (rc || !ctx->packed_ready) { b2_close_packed(ctx); ctx->packed_failed = true; }
    return rc;
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
    if (mode >= 2) { B2_TRY(open_peers(ctx)); if (!ctx->peers_open) mode = 1; }
    if (mode == 3) { B2_TRY(open_packed(ctx)); if (!ctx->packed_ready) mode = 2; }
    ctx->exchange_mode = mode;
    cudaStream_t xs = overlap ? ctx->comm_stream : s;
    // Slab order: own block first, then the other slabs by the place of their first block in the pull order (rank+1,
    // rank+2, ...); a slab may start once the LAST of its blocks in that order has landed (wait_k).
    int wait_k[64] = {0};
    if (overlap) {
        int pos = 0;
        for (int sl = ctx->own_lo; sl < ctx->own_hi; sl++) ctx->h_order[pos++] = sl;
        std::vector<std::pair<int, int>> rest;  // (first place in the pull order, slab)
        for (int sl = 0; sl < ctx->slabs; sl++) {
            if (sl >= ctx->own_lo && sl < ctx->own_hi) continue;
            int first = W, last = 0;
            for (int b = ctx->slab_blk_lo[sl]; b < ctx->slab_blk_hi[sl]; b++) {
                const int k = (b - ctx->rank + W) % W;
                first = std::min(first, k);
                last = std::max(last, k);
            }
            rest.push_back(std::make_pair(first, sl));
            wait_k[sl] = last;
        }
        std::sort(rest.begin(), rest.end());
        for (auto &p : rest) ctx->h_order[pos++] = p.second;
        for (int k = 0; k < W; k++)
            if (!ctx->ev_upto[k]) B2_CHECK(ctx, cudaEventCreateWithFlags(&ctx->ev_upto[k], cudaEventDisableTiming));
    }
    uint64_t *barrier_word = (uint64_t *)(ctx->d_norm + 4);
    bool in_flight = false, packed_in_flight = false;
    B2_CHECK(ctx, cudaEventRecord(ctx->ev0, s));
    for (int it = 0; it < steps; it++) {
        if (overlap) {
            for (int pos = 0; pos < ctx->slabs; pos++) ctx->slab_wait[pos] = nullptr;
            if (in_flight && mode == 1) ctx->slab_wait[per_block] = ctx->ev_xdone;
            if (in_flight && mode >= 2)
                for (int pos = per_block; pos < ctx->slabs; pos++) ctx->slab_wait[pos] = ctx->ev_upto[wait_k[ctx->h_order[pos]]];
            ctx->order_n = ctx->slabs;
        }
        // exchange 3: the previous step left the other blocks PACKED (only the columns this rank gathers); the whole
        // x is there before the first step and again after the last (which pulls whole blocks)
        ctx->packed_active = packed_in_flight;
        const int rc = b2_spmv_launch(ctx, spmv_flags, s);
        ctx->order_n = 0;
        ctx->packed_active = false;
        if (rc) return rc;
        if (!is_u64) {
            sumsq_partial_kernel<<<RED_BLOCKS, RED_THREADS, 0, s>>>((const double *)yb, ctx->local_rows, partial);
            sumsq_final_kernel<<<1, RED_THREADS, 0, s>>>(partial, RED_BLOCKS, norm2);
            if (ctx->world > 1) NCCL_TRY(ctx, g_nccl.allreduce(norm2, norm2, 1, NCCL_FLOAT64, NCCL_SUM, ctx->comm, s));
        }
        // peer pulls, u64 (no norm to agree on): nobody overwrites its block before everybody has pulled it
        if (mode >= 2 && is_u64) NCCL_TRY(ctx, g_nccl.allreduce(barrier_word + 1, barrier_word + 1, 1, NCCL_UINT64, NCCL_SUM, ctx->comm, s));
        scale_into_x_kernel<<<ctx->sms * 8, 256, 0, s>>>(yb, ctx->local_rows, norm2, is_u64, xb);
        ctx->launches += is_u64 ? 1 : 3;
        const bool packed_now = mode == 3 && it < steps - 1;
        if (packed_now) {  // my block, packed once per peer (the buffers the peers pull from)
            pack_kernel<<<ctx->sms * 8, 256, 0, s>>>(xb, ctx->local_rows, W, ctx->rank, ctx->d_pack_bits, ctx->d_pack_pref, ctx->pack_words,
                                                      ctx->d_pack, ctx->d_pack_off);
            B2_CHECK(ctx, cudaGetLastError());
            ctx->launches += 1;
        }
        // exchange time of a representative step -> stat "xchgtime_ns" (packed: the last packed one, else the last)
        const bool timed = ctx->world > 1 && (mode == 3 && steps > 1 ? it == steps - 2 : it == steps - 1);
        if (overlap) {
            B2_CHECK(ctx, cudaEventRecord(ctx->ev_xready, s));
            B2_CHECK(ctx, cudaStreamWaitEvent(xs, ctx->ev_xready, 0));
        }
        if (timed) {
            if (!ctx->ev_x0) { B2_CHECK(ctx, cudaEventCreate(&ctx->ev_x0)); B2_CHECK(ctx, cudaEventCreate(&ctx->ev_x1)); }
            B2_CHECK(ctx, cudaEventRecord(ctx->ev_x0, xs));
        }
        if (mode >= 2) {
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
                if (packed_now) {
                    if (ctx->need_cnt[r])
                        B2_CHECK(ctx, cudaMemcpyAsync(ctx->d_unpack + ctx->unpack_off[r], (const uint64_t *)ctx->pack_peer[r] + ctx->peer_pack_off[r],
                                                      ctx->need_cnt[r] * 8, cudaMemcpyDefault, ps));
                } else if (e > b)
                    B2_CHECK(ctx, cudaMemcpyAsync((uint64_t *)ctx->d_x + b, (const uint64_t *)ctx->x_peer[r] + b, (e - b) * 8,
                                                  cudaMemcpyDefault, ps));
                B2_CHECK(ctx, cudaEventRecord(ctx->ev_blk[r], ps));
                B2_CHECK(ctx, cudaStreamWaitEvent(xs, ctx->ev_blk[r], 0));
                B2_CHECK(ctx, cudaEventRecord(ctx->ev_upto[k], xs));  // blocks rank+1 .. rank+k are all here
            }
        } else {
            B2_TRY(allgather_x(ctx, xs));
        }
        if (timed) B2_CHECK(ctx, cudaEventRecord(ctx->ev_x1, xs));
        if (overlap) {
            B2_CHECK(ctx, cudaEventRecord(ctx->ev_xdone, xs));
            in_flight = true;
            packed_in_flight = packed_now;
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
