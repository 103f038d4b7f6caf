// common.cuh -- context, error plumbing and PTX helpers shared by the kernels.
#ifndef B200SPMV_COMMON_CUH_
#define B200SPMV_COMMON_CUH_

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>

#include "../../include/b200spmv.h"

struct ncclComm;
constexpr int Y_CHUNKS = 16;  // b200spmv_exec: row chunks of the y upload / accumulate / download pipeline

// ---------------------------------------------------------------------------------------
// Device-resident matrix, one per context.
//   CSC (as the reference stores it: software/SparseMatrix.h:60-69) is kept only until the
//   CSR is built unless keep_csc is set (tests / get_csc).
//   CSR is stored per column slab s: rowptr_s[local_rows+1] (offsets into the rank-local
//   colind/vals arrays), so that the x slice gathered by one slab pass fits the L2 set-aside.
// ---------------------------------------------------------------------------------------
struct DeviceCsr {
    uint32_t *rowptr = nullptr;  // [slabs][local_rows+1], absolute offsets into colind/vals
    uint32_t *colind = nullptr;  // [local_nnz (+pad)], GLOBAL column ids
    void *vals = nullptr;        // [local_nnz (+pad)] 8-byte
    // flagged format (kernels 3 and 6): bit 31 of colind marks the first entry of a row; seg_row maps
    // the n-th non-empty row of a slab to its row id; tile_seg[t] = flags before entry t*FLAG_TILE
    uint32_t *seg_row = nullptr;    // [slabs][rowptr_stride]  (nseg+1 used, last = NO_ROW)
    uint32_t *tile_seg = nullptr;   // [slabs] tables, slab s at h_ftile_base[s] + s, ntiles_s + 1 entries
    uint32_t *warp_seg = nullptr;   // same per 256-entry warp step (kernel 6), slab s at 8*h_ftile_base[s] + s
    uint32_t *carry_row = nullptr;  // [8 * tiles] row of a tile's (kernel 3) / chunk's (kernel 6) carry-out
    void *carry_val = nullptr;      // ... and the 8-byte partial of that row
};

struct b200spmv_ctx {
    int device = 0;
    int sms = 0;
    size_t l2_bytes = 0;
    size_t max_persist_bytes = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    int cuda_error = 0;

    // options
    int opt_slabs = 0;            // 0 = auto
    int64_t opt_persist_bytes = 0;  // 0 = device max
    int opt_persist = -1;         // L2 access-policy window + evict_last on x: -1 = as measured at setup (persist_tuned), 0 = off,
                                  // 1 = on; flag B200SPMV_NO_PERSIST switches it off for one call whatever this says
    int opt_require_aligned = 0;
    int opt_keep_csc = 0;
    int opt_kernel = 3;           // 3 = flagged segmented sum, one tile per CTA (default); 6 = the same by persistent warps (spmv.cu)
    int opt_stream_db = 0;        // kernel 6: register double buffer for the entry streams
    int opt_stream_pf = 1;        // kernel 6: L2 bulk prefetch (UBLKPF) distance in warp steps (0 = off)
    int opt_stream_occ = 0;       // kernel 6: CTAs per SM (0 = what the launch bounds allow: 2 with the double buffer, else 3)
    int opt_stream_vchunks = 1;   // kernel 6: chunks per warp (1 = one contiguous chunk per warp)
    int opt_dcsr_x16 = -1;        // which slabs run spmv_dcsr_kernel: -1 = measured per slab at setup (b2_tune_slab_kernels),
                                  // 0 = never, n > 0 = those with < n/16 entries per non-empty row
    int opt_prefetch_tiles = 148; // kernel 3: L2 bulk prefetch (UBLKPF) of the tile this many CTAs ahead (0 = off)
    int opt_own_slabs = 0;        // align the slabs with the ranks' blocks of x (set before setup; see own_lo): 1 = every block its own
                                  // slabs; 2 = the own block, and ONE slab each for all the blocks left and right of it
    int opt_exchange = 2;         // with own_slabs: 1 = NCCL all-gather under the own-block slabs, 2 = peer pulls, 3 = packed peer pulls
    int opt_pull_streams = 1;     // copy streams the peer pulls are spread over
    int opt_dcsr_maxrow = 64;     // ... and only when no row piece of the slab is longer than this
    int opt_pin_host = 0;         // see reg_ptr
    int opt_prefetch_x = 0;       // 1: stream each slab's slice of x into L2 (UBLKPF) before its kernel
    int opt_graph = 1;            // b200spmv_exec_device replays a captured CUDA graph of one step
    int opt_profile = 0;          // 1: bracket every merge-kernel launch with events (bench roofline)
    int opt_exec_gather = 1;      // b200spmv_exec with a communicator: upload 1/world of x, all-gather the rest over NVLink

    // matrix
    bool ready = false;
    int dtype = B200SPMV_F64;
    uint32_t rows = 0, cols = 0;
    uint64_t nnz = 0;             // global
    uint32_t row_begin = 0, row_end = 0;  // this rank's row block
    uint32_t local_rows = 0;
    size_t rowptr_stride = 0;     // elements between the slabs' row-pointer arrays (multiple of 4)
    uint64_t local_nnz = 0;
    uint64_t nonempty_rows = 0;   // of this rank's block
    uint64_t upper_half_nnz = 0;  // entries of the block in columns >= cols/2 (column skew, choose_slabs)
    uint64_t band_nnz = 0;        // entries of the block within L2/16 worth of columns of the diagonal (choose_slabs)
    int slabs = 1;
    uint32_t slab_cols = 0;       // columns per slab (last may be short) when the slabs are uniform
    uint32_t h_slab_col[65] = {0};     // column boundaries of the slabs
    // iterated SpMV over several ranks (option own_slabs): slabs [own_lo, own_hi) cover exactly this rank's own
    // block of x.  b200spmv_power_iterate then hands b2_spmv_launch an explicit slab order (own block first, the
    // other blocks in the order they arrive) and, per position, an event to wait on before that slab starts.
    int own_lo = 0, own_hi = 0;
    int slab_blk_lo[64] = {0}, slab_blk_hi[64] = {0};  // with own_slabs: the ranks' blocks [lo, hi) slab s lies in
    cudaEvent_t ev_upto[64] = {nullptr};    // the first k blocks of the pull order have landed
    uint32_t *d_splits = nullptr;      // h_row_splits on the device (packed exchange)
    int order_n = 0;                   // == slabs while an order is in force, else 0
    int h_order[64] = {0};             // slab run at position p
    cudaEvent_t slab_wait[64] = {nullptr};  // by position; nullptr = no wait
    cudaEvent_t ev_xready = nullptr, ev_xdone = nullptr, ev_bar = nullptr;
    cudaEvent_t ev_blk[64] = {nullptr};     // block r of x has landed (peer pulls)
    cudaStream_t comm_stream = nullptr;
    cudaStream_t pull_stream[4] = {nullptr};
    void *x_peer[64] = {nullptr};      // the other ranks' d_x, mapped through CUDA IPC
    bool peers_open = false, peers_failed = false;
    int exchange_mode = 0;             // what the last power iteration did: 0 after the step (NCCL), 1 all-gather
                                       // under the own-block slabs, 2 peer pulls block by block, 3 packed peer pulls
    // exchange 3, packed blocks (multi.cu open_packed): a rank gathers only from the columns its rows reference, so
    // the owner of a block packs x[those columns] per peer and the peer pulls the packed block; the remote slabs
    // then index the packed block through a renumbered copy of colind
    bool packed_ready = false, packed_failed = false;
    bool packed_active = false;        // this launch: remote slabs read colind_packed / d_unpack
    uint32_t *colind_packed = nullptr; // csr.colind with the remote slabs' columns renumbered (flags kept)
    uint64_t *d_unpack = nullptr;      // the packed blocks of the other ranks as pulled: block r at unpack_off[r]
    uint64_t unpack_off[65] = {0}, need_cnt[64] = {0};  // (colind_packed holds positions in the whole of d_unpack)
    uint64_t *d_unpack_off = nullptr;
    uint64_t *d_pack = nullptr;        // my block packed for each peer: for rank r at pack_off[r], pack_cnt[r] values
    uint64_t pack_off[65] = {0}, pack_cnt[64] = {0};
    uint64_t *d_pack_off = nullptr;    // pack_off on the device (pack kernel argument)
    uint32_t *d_pack_bits = nullptr;   // [world][pack_words]: bit c of row r = rank r needs column own_begin + c
    uint32_t *d_pack_pref = nullptr;   // exclusive prefix of its popcounts
    uint32_t pack_words = 0;
    void *pack_peer[64] = {nullptr};   // the other ranks' d_pack (CUDA IPC)
    uint64_t peer_pack_off[64] = {0};  // where MY packed block starts inside rank r's d_pack
    uint32_t h_ftile_base[65] = {0};   // flagged-kernel tiles per slab (prefix)
    uint64_t h_slab_off[65] = {0};     // entry offset of each slab in colind/vals (padded to FLAG_TILE)
    uint32_t h_slab_nnz[64] = {0};
    uint32_t h_nseg[64] = {0};         // non-empty rows per slab
    uint32_t h_maxrow[64] = {0};       // longest row piece per slab
    int8_t h_slab_dcsr[64] = {0};      // measured choice per slab (opt_dcsr_x16 < 0): 1 = thread-per-row kernel
    uint64_t tune_mask = 0;            // b2_tune_slab_kernels: the slabs being measured ...
    int tune_force = 0;                // ... 1 = with the flagged kernel, 2 = with the thread-per-row kernel
    uint64_t tunetime_ns = 0;
    bool flagged = false;              // colind carries row-start flags in bit 31
    uint32_t colmask = 0xFFFFFFFFu;    // mask that strips the flag
    uint64_t padded_nnz = 0;
    uint32_t total_tiles = 0;          // 2048-entry tiles over all slabs
    DeviceCsr csr;
    // CSC copy (rank-local entries, global ids)
    uint32_t *csc_colptr = nullptr, *csc_rowind = nullptr;
    void *csc_vals = nullptr;

    // vectors (device): x is the full replicated vector, y the full-length result (only
    // [row_begin,row_end) is written by this rank)
    void *d_x = nullptr, *d_y = nullptr;
    void *d_yin = nullptr;        // the caller's y while t = A*x is computed (b200spmv_exec, accumulate)
    // b200spmv_exec, result download: the kernel of the LAST slab is launched in Y_CHUNKS pieces of consecutive
    // tiles; the rows below h_chunk_row[k + 1] are final once piece k and its fix-up are done (ev_rowchunk[k]),
    // so their download (other PCIe direction) runs under the remaining pieces instead of after the whole step
    uint32_t h_chunk_tile[Y_CHUNKS + 1] = {0}, h_chunk_row[Y_CHUNKS + 1] = {0};
    bool chunks_ready = false;         // boundaries valid (flagged layout, last slab)
    bool want_row_chunks = false;      // set by b200spmv_exec around the launch
    bool did_row_chunks = false;       // ... and answered by b2_spmv_launch
    cudaEvent_t ev_rowchunk[Y_CHUNKS] = {nullptr};
    cudaEvent_t ev_tl[5] = {nullptr};  // exec timeline: start, x up, caller's y up, kernels done, download done
    uint64_t tl_ns[4] = {0};           // the last four, relative to the first (stats exec_*_ns)
    cudaStream_t d2h_stream = nullptr;
    cudaEvent_t ev_ychunk[Y_CHUNKS] = {nullptr};  // chunk k of the caller's y has landed in d_yin
    cudaEvent_t ev_slab[64] = {nullptr};  // slice s of x has landed (b200spmv_exec uploads x slab by slab)
    bool wait_slab_events = false;
    cudaEvent_t ev_x0 = nullptr, ev_x1 = nullptr;  // around the x exchange of the last power-iteration step
    uint64_t exec_h2d_bytes = 0, exec_d2h_bytes = 0;  // what the last b200spmv_exec moved over PCIe on this rank
    uint64_t xchgtime_ns = 0;
    // option pin_host: the caller's pageable x / y are page-locked in place (cudaHostRegister) the first time they
    // are seen and stay so until another pointer takes the slot or the context goes: SpMV objects keep their
    // vectors for life (software/SpMV.h:31-33), so every exec after the first copies at pinned speed
    void *reg_ptr[2] = {nullptr, nullptr};
    size_t reg_bytes[2] = {0, 0};
    bool persist_set = false;
    void *noalloc_scratch = nullptr;   // device address of spmv.cu's static carry scratch on this context's GPU (b200spmv_exec_noalloc)
    bool persist_tuned = true;    // what b2_tune_slab_kernels measured (window on unless off is > 1 % faster)
    int tune_persist = -1;        // during that measurement: 0 / 1 = force

    // multi-GPU
    ncclComm *comm = nullptr;
    int rank = 0, world = 1;
    uint32_t *h_row_splits = nullptr;  // [world+1]
    double *d_norm = nullptr;          // scratch for power iteration

    // stats
    uint64_t spmvtime_ns = 0, xfertime_ns = 0, setuptime_ns = 0, transposetime_ns = 0;
    uint64_t launches = 0;
    uint64_t launches_per_step = 0;
    cudaEvent_t prof_ev[128] = {nullptr};
    int prof_n = 0;               // events recorded since the last read
    uint64_t merge_kernel_ns = 0, merge_kernel_launches = 0;
    uint64_t prof_slab_ns[64] = {0};  // the same, by position of the launch within a step
    int last_mode = 0;
};

#define B2_CHECK(ctx, call)                                                              \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess) {                                                          \
            char buf_[512];                                                               \
            snprintf(buf_, sizeof buf_, "%s:%d: %s -> %s", __FILE__, __LINE__, #call,     \
                     cudaGetErrorString(e_));                                             \
            (ctx)->err = buf_;                                                            \
            (ctx)->cuda_error = (int)e_;                                                  \
            return e_ == cudaErrorMemoryAllocation ? B200SPMV_ENOMEM : B200SPMV_ECUDA;    \
        }                                                                                 \
    } while (0)

#define B2_TRY(expr)             \
    do {                         \
        int rc_ = (expr);        \
        if (rc_ != 0) return rc_; \
    } while (0)

static inline int b2_fail(b200spmv_ctx *ctx, int code, const char *msg)
{
    if (ctx) ctx->err = msg;
    return code;
}

template <typename T>
static inline int b2_alloc(b200spmv_ctx *ctx, T **p, size_t count)
{
    *p = nullptr;
    // pad every array by 256 bytes so 16-byte-granular tile loads may overrun the logical end
    B2_CHECK(ctx, cudaMalloc((void **)p, count * sizeof(T) + 256));
    return 0;
}
static inline int b2_alloc_bytes(b200spmv_ctx *ctx, void **p, size_t bytes)
{
    *p = nullptr;
    B2_CHECK(ctx, cudaMalloc(p, bytes + 256));
    return 0;
}
template <typename T>
static inline void b2_free(T *&p)
{
    if (p) cudaFree((void *)p);
    p = nullptr;
}

static inline unsigned b2_div_up(uint64_t a, uint64_t b) { return (unsigned)((a + b - 1) / b); }

// ---- module entry points (host side) ---------------------------------------------------
// scan.cu
int b2_exclusive_scan_u32(b200spmv_ctx *ctx, uint32_t *data, uint64_t n, cudaStream_t s);
size_t b2_scan_scratch_bytes(uint64_t n);
// sort.cu : stable LSD radix sort of (key, payload) pairs; key_bits significant bits
int b2_sort_pairs(b200spmv_ctx *ctx, uint32_t *keys, uint32_t *payload, uint32_t *keys_tmp,
                  uint32_t *payload_tmp, uint64_t n, int key_bits, bool payload_is_iota,
                  uint32_t **sorted_payload, cudaStream_t s);
// transpose.cu
int b2_compress(b200spmv_ctx *ctx, uint32_t nkeys, uint64_t n, const uint32_t *key, uint32_t key_mask,
                uint32_t key_sub, const uint32_t *src_ptr, uint32_t src_major, uint32_t src_pos_off,
                const uint32_t *src_other, const void *src_vals, uint32_t *dst_ptr, uint32_t dst_ptr_add,
                uint32_t *dst_ind, void *dst_vals, cudaStream_t s);
int b2_add_u32(b200spmv_ctx *ctx, uint32_t *a, uint64_t n, uint32_t add, cudaStream_t s);
int b2_count_band(b200spmv_ctx *ctx, const uint32_t *colptr, const uint32_t *rowind, uint32_t mask, uint32_t cols, uint32_t rows,
                  uint32_t window, uint64_t *out, cudaStream_t s);
int b2_validate_indices(b200spmv_ctx *ctx, const uint32_t *d_ind, uint64_t n, uint32_t mask, uint32_t limit,
                        const uint32_t *d_ptr, uint64_t nptr, cudaStream_t s);
int b2_count_nonempty(b200spmv_ctx *ctx, uint32_t nkeys, uint64_t n, const uint32_t *key, uint32_t key_mask,
                      uint32_t key_sub, uint64_t *out, cudaStream_t s);
// spmv.cu
int b2_build_flags(b200spmv_ctx *ctx);
constexpr uint32_t B2_FLAG_TILE = 2048;  // entries per CTA of kernel 3 (8 warp steps of kernel 6); slabs are padded to it
int b2_spmv_launch(b200spmv_ctx *ctx, int flags, cudaStream_t s);
int b2_tune_slab_kernels(b200spmv_ctx *ctx);
int b2_spmv_prepare(b200spmv_ctx *ctx, int flags, cudaStream_t s);
bool b2_last_slab_chunkable(const b200spmv_ctx *ctx, int flags);
void b2_profile_collect(b200spmv_ctx *ctx);
// synth.cu
int b2_synth_build_csc(b200spmv_ctx *ctx, int kind, uint32_t npr, uint64_t seed);
int b2_synth_fill_x(b200spmv_ctx *ctx, uint64_t seed);
// multi.cu
int b2_comm_destroy(b200spmv_ctx *ctx);
void b2_close_packed(b200spmv_ctx *ctx);
void b2_close_peers(b200spmv_ctx *ctx);
int b2_shard_csc(b200spmv_ctx *ctx, uint32_t index_mask);
void b2_exec_x_block(const b200spmv_ctx *ctx, int r, uint64_t *lo, uint64_t *hi);
int b2_exec_gather_x(b200spmv_ctx *ctx, cudaStream_t s);

// ---- device helpers --------------------------------------------------------------------
#if defined(__CUDACC__)

// L2 eviction policies.  The matrix streams are read once: evict_first, so they do not
// displace the gathered vector (the "vector cache" contents), which is evict_last.
// (On sm_100a the direct .L2::evict_* qualifiers exist only for 256-bit loads; narrower
// loads take a createpolicy descriptor through .L2::cache_hint.)
__device__ __forceinline__ uint64_t policy_evict_first()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_last()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
// streaming 128-bit load: no L1 allocation, L2 policy `pol`
__device__ __forceinline__ uint4 ld_stream_v4(const void *p, uint64_t pol)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p), "l"(pol));
    return r;
}
// streaming 256-bit load (Blackwell): four 8-byte values, 32-byte aligned
__device__ __forceinline__ void ld_stream_v4_b64(const void *p, uint64_t (&v)[4])
{
    asm volatile("ld.global.nc.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3])
                 : "l"(p));
}
__device__ __forceinline__ uint32_t ld_stream_u32(const void *p, uint64_t pol)
{
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(r) : "l"(p), "l"(pol));
    return r;
}
// gathered-vector load: read-only path, keep in L2
__device__ __forceinline__ uint64_t ld_gather_u64(const void *p, uint64_t pol)
{
    uint64_t r;
    asm volatile("ld.global.nc.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(r) : "l"(p), "l"(pol));
    return r;
}

// Row-block split of `rows` rows over `world` ranks, one thread block of >= world+1 threads:
// nnz-balanced -- split k = first row whose prefix of stored entries reaches k*nnz/world -- unless plain
// equal row counts are already balanced to 2 % (banded and uniform matrices): equal blocks let the iterated
// SpMV refresh x with one in-place ncclAllGather instead of a group of broadcasts.
__device__ __forceinline__ void b2_row_splits(const uint32_t *__restrict__ rowptr, uint32_t rows, int world,
                                              uint32_t *__restrict__ splits)
{
    const int k = threadIdx.x;
    if (k <= world) {
        uint32_t v;
        if (k == 0) v = 0;
        else if (k == world) v = rows;
        else {
            const uint64_t target = (uint64_t)rowptr[rows] * k / world;
            uint32_t lo = 0, hi = rows;
            while (lo < hi) {
                const uint32_t mid = lo + ((hi - lo) >> 1);
                if (rowptr[mid] < target) lo = mid + 1; else hi = mid;
            }
            v = lo;
        }
        splits[k] = v;
    }
    __syncthreads();
    if (k == 0 && world > 1 && rows % (uint32_t)world == 0) {
        const uint32_t per = rows / (uint32_t)world;
        uint64_t worst = 0;
        for (int r = 0; r < world; r++) {
            const uint64_t n = (uint64_t)rowptr[(r + 1) * per] - rowptr[r * per];
            worst = n > worst ? n : worst;
        }
        if (worst * 100ull * (uint64_t)world <= 102ull * (uint64_t)rowptr[rows] + 100ull * (uint64_t)world)
            for (int r = 0; r <= world; r++) splits[r] = r * per;
    }
}

// semiring (chisel/frontend/SemiringOp.scala): fp64 = separate IEEE multiply and add, never
// contracted to FMA; u64 = wrapping.
template <int DT> struct Semi;
template <> struct Semi<B200SPMV_F64> {
    __device__ __forceinline__ static uint64_t mul(uint64_t a, uint64_t b)
    {
        return (uint64_t)__double_as_longlong(__dmul_rn(__longlong_as_double((long long)a),
                                                         __longlong_as_double((long long)b)));
    }
    __device__ __forceinline__ static uint64_t add(uint64_t a, uint64_t b)
    {
        return (uint64_t)__double_as_longlong(__dadd_rn(__longlong_as_double((long long)a),
                                                         __longlong_as_double((long long)b)));
    }
};
template <> struct Semi<B200SPMV_U64> {
    __device__ __forceinline__ static uint64_t mul(uint64_t a, uint64_t b) { return a * b; }
    __device__ __forceinline__ static uint64_t add(uint64_t a, uint64_t b) { return a + b; }
};
#endif  // __CUDACC__

#endif
