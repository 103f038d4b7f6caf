/*
 * b200spmv.h -- C ABI of the B200-native SpMV engine (libb200spmv.so).
 *
 * This is the drop-in boundary for ONE path of maltanar/spmv-vector-cache: the
 * column-streaming y += A*x of software/SoftwareSpMV.cpp:50-70, as every SpMV subclass
 * exposes it (software/SpMV.h:8-35: ctor(A, x, y), exec(), statInt(), statKeys()).
 * A C++ subclass `CudaSpMV : public SpMV` (spmv-vector-cache_b200/host/CudaSpMV.h) binds
 * these entry points; INTEGRATION.md shows the lines a maintainer adds to the reference.
 *
 * Conventions: plain pointers and sizes only; every call returns 0 on success or a
 * negative B200SPMV_E* code (never throws); b200spmv_last_error() gives the text.
 * A context is bound to one GPU and is not thread-safe (the reference is a single
 * bare-metal thread, software/main.cpp).  There is NO CPU fallback: without a GPU
 * b200spmv_create fails with B200SPMV_ECUDA.
 */
#ifndef B200SPMV_H_
#define B200SPMV_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200SPMV_API __attribute__((visibility("default")))
#else
#define B200SPMV_API
#endif

typedef struct b200spmv_ctx b200spmv_ctx;

/* element type of values, x and y: SpMVData (software/SparseMatrix.h:6 is `double`; the
 * *-uint64 matrix directories use the wrapping integer semiring of
 * chisel/frontend/SemiringOp.scala:34-44). */
#define B200SPMV_F64 0
#define B200SPMV_U64 1

/* exec flags */
#define B200SPMV_ACCUMULATE 1   /* y += A*x (SoftwareSpMV.cpp:62).  Without it: y = A*x, what the
                                   FPGA variants compute after zero-filling (NewCache init phase) */
#define B200SPMV_STRICT_ORDER 2 /* row-sequential kernel, separate mul+add in the reference's
                                   summation order: fp64 results memcmp-equal to SoftwareSpMV */
#define B200SPMV_NO_PERSIST 4   /* do not pin x in L2 (A/B switch for the "vector cache") */

/* synthetic matrix kinds (BASELINE.json configs 3-5) */
#define B200SPMV_SYNTH_UNIFORM 0
#define B200SPMV_SYNTH_RMAT 1
#define B200SPMV_SYNTH_BANDED 2
#define B200SPMV_SYNTH_BAND 3 /* pure band of nnz_per_row consecutive columns (locality / HBM-roofline probe) */

/* error codes */
#define B200SPMV_OK 0
#define B200SPMV_EINVAL (-1)   /* bad argument (zero dimension: SparseMatrix.cpp:33-38) */
#define B200SPMV_ECUDA (-2)    /* CUDA runtime error, or no GPU */
#define B200SPMV_ENCCL (-3)
#define B200SPMV_ESTATE (-4)   /* call out of order (exec before setup, ...) */
#define B200SPMV_ENOMEM (-5)
#define B200SPMV_EALIGN (-6)   /* host array not 64-byte aligned (HardwareSpMV.cpp:6-16) */

/* ---- lifetime ---------------------------------------------------------------------- */
/* Replaces HWSpMVFactory::make's hardware probe (software/HWSpMVFactory.cpp:20-38). */
B200SPMV_API int b200spmv_create(b200spmv_ctx **out, int device);
B200SPMV_API void b200spmv_destroy(b200spmv_ctx *ctx);
B200SPMV_API const char *b200spmv_last_error(const b200spmv_ctx *ctx); /* ctx may be NULL: last create error */

/* Tunables, set before setup: "slabs" (column slabs, 0 = auto: sized so that a slab's slice of x fits L2, capped by the
 * entries per row and slab, 1 for a matrix whose entries sit within a narrow band around the diagonal -- stat "band_nz"), "persist_bytes"
 * (L2 set-aside cap, 0 = device max), "persist" (the L2 access-policy window + evict_last hint on x: -1, the default =
 * kept unless setup measured that dropping it is more than 1 % faster; 0 = off; 1 = on; stat "persist" tells which is
 * in force; flag B200SPMV_NO_PERSIST drops it for one call), "require_aligned" (1 = enforce the 64-byte rule),
 * "keep_csc" (1 = keep the device CSC copy for b200spmv_get_csc), "profile", "kernel" (3 = flagged
 * segmented sum with one tile per CTA, the default; 6 = the same by persistent warps, with its tuning knobs
 * "stream_db", "stream_pf", "stream_occ", "stream_vchunks"), "world" then "rank" (take the
 * place of rank r of w in the nnz-balanced row-block split WITHOUT a communicator: a single SpMV
 * needs no collective; b200spmv_power_iterate does and refuses).
 * Kernel selection per slab: "dcsr_x16" (which slabs run the thread-per-row kernel: -1, the default = whichever
 * kernel was faster when setup timed both on the slab; 0 = never; n > 0 = slabs with fewer than n/16 entries per
 * non-empty row) and "dcsr_maxrow" (... and only slabs with no row piece longer than this, default 64; stats
 * "slab_dcsr_<i>", "tunetime_ns"); "prefetch_tiles" (kernel 3: L2 bulk prefetch distance in CTAs, default 148), "prefetch_x" (1 = stream a
 * slab's slice of x into L2 before its kernel; measured useless, default 0), "graph" (b200spmv_exec_device
 * replays a CUDA graph of one step, default 1).
 * "pin_host" (1 = b200spmv_exec page-locks the caller's pageable x / y in place with cudaHostRegister the first
 * time it sees them, vectors >= 256 KB, and keeps them so until another pointer is passed or the context is
 * destroyed -- for callers whose vectors outlive the context's use of them, as SpMV objects' do; default 0).
 * For iterated SpMV over several ranks: "own_slabs" (1 = cut the column slabs along the ranks' blocks of
 * x, so that b200spmv_power_iterate can run the slabs of a rank's own block while the other blocks are
 * still on the wire; 2 = the same for the own block, but ONE slab for all the blocks left of it and one for
 * all the blocks right of it: fewer passes over y), "exchange" (with own_slabs: 2 = the other blocks are pulled
 * from the peers' x over NVLink by the copy engines through CUDA IPC mappings, block by block, each slab starting
 * when the last of its blocks has landed -- the default, falls back to 1 where peers cannot be mapped; 3 = the same
 * with PACKED blocks: the owner packs, per peer, only the columns that peer's rows gather, the peer pulls the packed
 * block and its remote slabs index it through a renumbered copy of the column indices; x is whole again when
 * b200spmv_power_iterate returns; 1 = one NCCL all-gather under the own-block slabs), "pull_streams" (copy streams of
 * exchanges 2 and 3, default 1).  Stat "exchange_mode" reports what the last b200spmv_power_iterate did (0 = NCCL
 * all-gather after the step). */
B200SPMV_API int b200spmv_set_option(b200spmv_ctx *ctx, const char *key, int64_t value);

/* ---- setup: upload CSC, transpose to CSR on the device (once, amortised) ------------- */
/* Replaces HardwareSpMV*::setupRegs+init (software/HardwareSpMVNewCache.cpp:31-55) and the
 * loader contract of SparseMatrix (SparseMatrix.h:18-70): colptr u32[cols+1], rowind
 * u32[nnz], vals 8-byte[nnz].  index_mask strips row-marker bits (0x3FFFFFFF when the
 * matrix went through SparseMatrix::markRowStarts, SparseMatrix.cpp:64; else 0xFFFFFFFF).
 * With a communicator attached (b200spmv_comm_init) every rank passes the same arrays and
 * keeps its nnz-balanced row block. */
B200SPMV_API int b200spmv_setup_csc(b200spmv_ctx *ctx, uint32_t rows, uint32_t cols, uint32_t nnz,
                       const uint32_t *colptr, const uint32_t *rowind, const void *vals,
                       int dtype, uint32_t index_mask);

/* Same, but the matrix is generated on the device (counter-based, see csrc/synth_rng.h),
 * laid out as the reference's CSC, then transposed by the same kernels.  For RMAT,
 * rows = cols = 2^scale and nnz_per_row is the edge factor. */
B200SPMV_API int b200spmv_setup_synthetic(b200spmv_ctx *ctx, int kind, uint32_t rows, uint32_t cols,
                             uint32_t nnz_per_row, uint64_t seed, int dtype);

/* Cross-check of a synthetic u64 matrix: recomputes A*x for this rank's rows straight from the generator's
 * definition (csrc/synth_rng.h) -- R-MAT: one wrapping atomic add per edge, the other kinds one thread per row; no
 * CSC, transpose, CSR or SpMV kernel involved -- with the x in HBM, and counts the rows where the y in HBM differs.
 * For the full-size parity tests of the configs too large for a host oracle. */
B200SPMV_API int b200spmv_check_synthetic_u64(b200spmv_ctx *ctx, int kind, uint32_t nnz_per_row, uint64_t seed,
                                 uint64_t *mismatches);

/* ---- exec: the SpMV::exec() contract (software/SpMV.h:12) ----------------------------- */
/* H2D x (and y when accumulating), kernels, D2H y into the caller's vector; blocking.
 * x_host: cols elements, y_host: rows elements (this rank's row block when sharded:
 * see b200spmv_stat_u64 "row_begin"/"row_end"; y_host still points at element 0 of the
 * full vector).
 * With a communicator attached (b200spmv_comm_init, world > 1) and x >= 1 MB this is a COLLECTIVE call: every rank
 * passes the same x, uploads only its 1/world of it over its own PCIe link and receives the rest from its peers
 * over NVLink (ncclAllGather); all ranks must call it together.  Option "exec_gather" = 0 restores the
 * every-rank-uploads-everything behaviour (no collective).  Stats exec_h2d_bytes / exec_d2h_bytes give what the
 * last call moved over PCIe on this rank. */
B200SPMV_API int b200spmv_exec(b200spmv_ctx *ctx, const void *x_host, void *y_host, int flags);

/* Device-resident variant for measurement: `iters` back-to-back SpMVs on the x / y already
 * in HBM (b200spmv_set_x / _set_y / _fill_x_synthetic).  ms_each[iters] (may be NULL)
 * receives CUDA-event times of each SpMV (all its kernels), on the launching stream. */
B200SPMV_API int b200spmv_exec_device(b200spmv_ctx *ctx, int iters, int flags, float *ms_each);

/* The SpMV for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * y = A*x (flags 0) or y += A*x (B200SPMV_ACCUMULATE) on DEVICE vectors the caller passes -- x_dev: cols elements,
 * y_dev: rows elements (a rank of a row-block split writes its own rows only) -- with the matrix the context was set
 * up with, kernel 3 (flagged segmented sum) on every column slab, launched on `stream` (a cudaStream_t; NULL = the
 * context's own stream, and then the call also waits for it).  The call makes no cudaMalloc / cudaFree and creates no
 * stream, event or graph: the per-tile carries, which setup normally allocates, live in a static device array inside
 * the library instead (3 MB per GPU that loads the library, held until the process exits).
 * Limit: every column slab of the matrix must have at most 2^18 tiles of 2048 entries (2^29 stored entries per slab);
 * larger matrices get B200SPMV_EINVAL with the slab and its tile count in b200spmv_last_error.  Other flags are
 * rejected (no strict order, no L2 window on the caller's x).
 * The scratch is shared by every call on the same GPU and keeps the last call's contents (nothing in it has to start
 * at zero: each tile's carry is written before it is read): calls whose kernels OVERLAP on one GPU -- different
 * streams, different contexts -- race; calls ordered on one stream do not.  fp64 results equal b200spmv_exec_device's
 * with kernel 3 bit for bit (same kernel, same tiles). */
B200SPMV_API int b200spmv_exec_noalloc(b200spmv_ctx *ctx, const void *x_dev, void *y_dev, int flags, void *stream);

B200SPMV_API int b200spmv_set_x(b200spmv_ctx *ctx, const void *x_host);          /* cols elements */
B200SPMV_API int b200spmv_set_y(b200spmv_ctx *ctx, const void *y_host);          /* rows elements (full) */
B200SPMV_API int b200spmv_fill_x_synthetic(b200spmv_ctx *ctx, uint64_t seed);    /* synth_x_bits(seed, i) */
B200SPMV_API int b200spmv_fill_y_zero(b200spmv_ctx *ctx);
B200SPMV_API int b200spmv_get_x(b200spmv_ctx *ctx, void *x_host);
B200SPMV_API int b200spmv_get_y(b200spmv_ctx *ctx, void *y_host);                /* this rank's rows only are written */

/* Read back what setup built (tests: stability of the transpose, oracle on synthetic data). */
B200SPMV_API int b200spmv_get_csr(b200spmv_ctx *ctx, uint32_t *rowptr, uint32_t *colind, void *vals);
B200SPMV_API int b200spmv_get_csc(b200spmv_ctx *ctx, uint32_t *colptr, uint32_t *rowind, void *vals);

/* ---- statistics: SpMV::statInt / statKeys (software/SpMV.h:28-29) ---------------------- */
/* keys: rows cols nz local_rows local_nz row_begin row_end spmvtime_ns xfertime_ns
 * setuptime_ns transposetime_ns xchgtime_ns (x exchange of the last power-iteration step) slabs l2_persist_bytes l2_bytes sms gpus rank mode
 * launches cuda_error tiles nonempty_rows (of this rank's block) exchange_mode; per slab i: slab_nnz_<i>,
 * slab_rows_<i> (non-empty rows), slab_maxrow_<i> (longest row piece); timeline of the last b200spmv_exec from
 * its first copy: exec_x_up_ns exec_y_up_ns exec_kernels_ns exec_done_ns; with option "profile"=1 also
 * merge_kernel_ns and merge_kernel_launches (CUDA-event time of the SpMV kernel alone, for the roofline) and
 * slab_ns_<i> (the same per slab) */
B200SPMV_API int b200spmv_stat_u64(b200spmv_ctx *ctx, const char *key, uint64_t *out);

/* ---- multi-GPU: one context (process) per GPU ------------------------------------------ */
/* 128-byte NCCL unique id made on rank 0 and distributed by the caller (MPI, torchrun, ...). */
B200SPMV_API int b200spmv_comm_unique_id(void *id128);
B200SPMV_API int b200spmv_comm_init(b200spmv_ctx *ctx, int rank, int world, const void *id128);

/* Power iteration (BASELINE.json configs[4]): per step y = A*x on the rank's row block,
 * global 2-norm, x <- y/||y|| refreshed on every rank (all-gather of the row blocks over
 * NVLink).  u64: x <- y, no normalisation.  x0 = 1/sqrt(n) (u64: 1) unless use_current_x.
 * ms_total = device time of all steps on this rank (caller takes the max over ranks). */
B200SPMV_API int b200spmv_power_iterate(b200spmv_ctx *ctx, int steps, int use_current_x, int flags,
                           float *ms_total, double *last_norm);

/* ---- row-marking preprocessing on the device (SparseMatrix.cpp:52-125) ------------------ */
/* Operate on host CSC index arrays in place, like the reference methods they replace. */
B200SPMV_API int b200spmv_mark_row_starts(b200spmv_ctx *ctx, uint32_t rows, uint32_t nnz, uint32_t *inds,
                             int reverse, int shift);
B200SPMV_API int b200spmv_max_alive(b200spmv_ctx *ctx, uint32_t rows, uint32_t nnz, uint32_t *inds,
                       uint32_t *out);
B200SPMV_API int b200spmv_max_col_span(b200spmv_ctx *ctx, uint32_t cols, uint32_t nnz, const uint32_t *colptr,
                          const uint32_t *inds, uint32_t *out);

/* ---- the vector cache as a model (chisel/cache-new/NoWMVectorCache.scala) ----------------------- */
/* What the FPGA's direct-mapped write-back cache over y (2^depth_log2 one-element lines) would do on the
 * CSC row-index stream `inds` (storage order): the counters software/HardwareSpMVNewCache.cpp:189-204
 * reports.  cms != 0: bit 31 of an index is the row-start mark of SparseMatrix::markRowStarts and such a
 * first touch does not read DRAM (cold-miss skip).
 * out[7] = reads, readMisses, conflictEvictions, coldMissSkips, ddrReads, flushWrites, hits. */
B200SPMV_API int b200spmv_vector_cache_model(b200spmv_ctx *ctx, uint32_t nnz, const uint32_t *inds, int depth_log2,
                                int cms, uint64_t *out);

#ifdef __cplusplus
}
#endif
#endif /* B200SPMV_H_ */
