/*
 * spmv_oracle.c -- CPU restatement of the spmv-vector-cache hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (include/, the
 * spmv-vector-cache_b200 package) may link, import or execute this file.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, and only as the checker or the CPU baseline.
 *
 * Parity: PINNED.  tests/test_oracle.py checks these functions against the five
 * golden.bin fixtures the reference ships (matrices/{circuit204,i64,i1k,i64k,row64k}),
 * against known-answer tests of chisel/tests/TestSpMVFrontend.scala:121-183, and
 * (in the build container, where /root/reference exists) against the reference's own
 * SoftwareSpMV.cpp / SparseMatrix.cpp / csr2csc.c compiled unmodified into
 * oracle/_ref/ (see oracle/Makefile).
 *
 * Every function cites the reference lines (relative to /root/reference) it restates.
 * Compile with -ffp-contract=off (no FMA fusion: the reference does a separate
 * multiply and add, software/SoftwareSpMV.cpp:62, chisel/frontend/SemiringOp.scala:94-144)
 * and never with -ffast-math / FTZ.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <time.h>

#define ORACLE_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------- */
/* y += A*x over CSC streams, columns ascending, entries in storage order.    */
/* software/SoftwareSpMV.cpp:59-64.  index_mask strips the row-start/row-end  */
/* marker bits (SparseMatrix.cpp:64 uses 0x3FFFFFFF); pass 0xFFFFFFFF for raw.*/
/* ------------------------------------------------------------------------- */
ORACLE_API void oracle_spmv_csc_f64(uint32_t rows, uint32_t cols,
                                    const uint32_t *colptr, const uint32_t *rowind,
                                    const double *vals, const double *x, double *y,
                                    uint32_t index_mask)
{
    (void)rows;
    for (uint32_t col = 0; col < cols; col++) {
        double xj = x[col];
        for (uint32_t e = colptr[col]; e < colptr[col + 1]; e++) {
            double prod = vals[e] * xj;            /* separate mul ...          */
            y[rowind[e] & index_mask] += prod;     /* ... then add (no FMA)     */
        }
    }
}

/* Same loop with SpMVData = unsigned 64-bit: wrapping mul and add (mod 2^64), */
/* the semantics of chisel/frontend/SemiringOp.scala:34-44.                    */
ORACLE_API void oracle_spmv_csc_u64(uint32_t rows, uint32_t cols,
                                    const uint32_t *colptr, const uint32_t *rowind,
                                    const uint64_t *vals, const uint64_t *x, uint64_t *y,
                                    uint32_t index_mask)
{
    (void)rows;
    for (uint32_t col = 0; col < cols; col++) {
        uint64_t xj = x[col];
        for (uint32_t e = colptr[col]; e < colptr[col + 1]; e++)
            y[rowind[e] & index_mask] += vals[e] * xj;
    }
}

/* Row-sequential CSR form of the same sum.  With a STABLE transpose this visits  */
/* each row's entries in exactly the order SoftwareSpMV.cpp:59-64 adds them, so it */
/* is bit-identical to oracle_spmv_csc_f64 -- the property strict mode relies on.  */
ORACLE_API void oracle_spmv_csr_f64(uint32_t rows, const uint32_t *rowptr,
                                    const uint32_t *colind, const double *vals,
                                    const double *x, double *y)
{
    for (uint32_t r = 0; r < rows; r++) {
        double acc = y[r];
        for (uint32_t p = rowptr[r]; p < rowptr[r + 1]; p++) {
            double prod = vals[p] * x[colind[p]];
            acc += prod;
        }
        y[r] = acc;
    }
}

ORACLE_API void oracle_spmv_csr_u64(uint32_t rows, const uint32_t *rowptr,
                                    const uint32_t *colind, const uint64_t *vals,
                                    const uint64_t *x, uint64_t *y)
{
    for (uint32_t r = 0; r < rows; r++) {
        uint64_t acc = y[r];
        for (uint32_t p = rowptr[r]; p < rowptr[r + 1]; p++)
            acc += vals[p] * x[colind[p]];
        y[r] = acc;
    }
}

/* ------------------------------------------------------------------------- */
/* Compressed-format transpose: the counting sort of software/csr2csc.c:11-39 */
/* (count :20, running sum :22, stable cursor scatter :27-33, shift back       */
/* :36-38), restated for unsigned 32-bit indices and opaque 8-byte values so   */
/* it serves both directions: CSC->CSR (src_major=cols, dst_major=rows) and    */
/* CSR->CSC (the direction csr2csc.c itself is written in).                    */
/*   src_ptr[src_major+1], src_ind[nnz] (ids in the dst dimension), src_val    */
/*   dst_ptr[dst_major+1], dst_ind[nnz] (ids in the src dimension), dst_val    */
/* vals may be NULL (pattern only, csr2csc.c "if (a)").                        */
/* ------------------------------------------------------------------------- */
ORACLE_API void oracle_transpose(uint32_t src_major, uint32_t dst_major, uint32_t nnz,
                                 const uint32_t *src_ptr, const uint32_t *src_ind,
                                 const uint64_t *src_val, uint32_t *dst_ptr,
                                 uint32_t *dst_ind, uint64_t *dst_val, uint32_t index_mask)
{
    for (uint32_t i = 0; i <= dst_major; i++) dst_ptr[i] = 0;
    for (uint32_t e = 0; e < nnz; e++) dst_ptr[(src_ind[e] & index_mask) + 1]++;
    for (uint32_t i = 0; i < dst_major; i++) dst_ptr[i + 1] += dst_ptr[i];
    for (uint32_t s = 0; s < src_major; s++) {
        for (uint32_t e = src_ptr[s]; e < src_ptr[s + 1]; e++) {
            uint32_t k = src_ind[e] & index_mask;
            uint32_t l = dst_ptr[k]++;
            dst_ind[l] = s;
            if (src_val) dst_val[l] = src_val[e];
        }
    }
    for (uint32_t i = dst_major; i > 0; i--) dst_ptr[i] = dst_ptr[i - 1];
    dst_ptr[0] = 0;
}

/* ------------------------------------------------------------------------- */
/* Row-marking preprocessing (cold-miss skip), software/SparseMatrix.cpp.      */
/* ------------------------------------------------------------------------- */

/* SparseMatrix::markRowStarts, SparseMatrix.cpp:52-90: flag the first (or, when */
/* reverse, the last) storage-order occurrence of every row id in bit `shift`.   */
ORACLE_API void oracle_mark_row_starts(uint32_t rows, uint32_t nnz, uint32_t *inds,
                                       int reverse, int shift)
{
    uint32_t enc_size = rows / 32 + 1;
    uint32_t *seen = (uint32_t *)calloc(enc_size, sizeof(uint32_t));
    for (uint32_t i = 0; i < nnz; i++) {
        uint32_t e = reverse ? (nnz - 1 - i) : i;
        uint32_t r = inds[e] & 0x3FFFFFFFu;
        uint32_t w = r / 32, b = r % 32;
        if ((seen[w] & (1u << b)) == 0) {
            seen[w] |= (1u << b);
            inds[e] |= (1u << shift);
        }
    }
    free(seen);
}

/* SparseMatrix::clearRowMarkings, SparseMatrix.cpp:121-125. */
ORACLE_API void oracle_clear_row_markings(uint32_t nnz, uint32_t *inds, uint32_t mask)
{
    for (uint32_t e = 0; e < nnz; e++) inds[e] &= mask;
}

/* SparseMatrix::maxAlive, SparseMatrix.cpp:92-108.  Marks bit 31 (first) and bit */
/* 30 (last) in place, like the reference, and leaves them set.  Note the         */
/* reference order inside the loop: +1 for a start, -1 for an end, THEN max, so a */
/* row with a single entry never counts as alive.                                 */
ORACLE_API uint32_t oracle_max_alive(uint32_t rows, uint32_t nnz, uint32_t *inds)
{
    oracle_mark_row_starts(rows, nnz, inds, 0, 31);
    oracle_mark_row_starts(rows, nnz, inds, 1, 30);
    uint32_t max_alive = 0, cur = 0;
    for (uint32_t e = 0; e < nnz; e++) {
        if (inds[e] & (1u << 31)) cur += 1;
        if (inds[e] & (1u << 30)) cur -= 1;
        if (cur > max_alive) max_alive = cur;
    }
    return max_alive;
}

/* SparseMatrix::maxColSpan, SparseMatrix.cpp:110-119: max over columns of        */
/* inds[colptr[c+1]-1] - inds[colptr[c]] in unsigned arithmetic on the raw words.  */
/* Like the reference, an EMPTY interior column is not skipped (it compares the    */
/* neighbouring columns' entries).  Only reads that would leave the array -- an    */
/* empty leading column (colptr[c+1]==0) or trailing one (colptr[c]==nnz) -- are   */
/* skipped, for memory safety; the reference's callers never hit them.             */
ORACLE_API uint32_t oracle_max_col_span(uint32_t cols, uint32_t nnz, const uint32_t *colptr,
                                        const uint32_t *inds)
{
    uint32_t best = 0;
    for (uint32_t c = 0; c < cols; c++) {
        if (colptr[c + 1] == 0 || colptr[c] >= nnz) continue; /* memory-safety guard only */
        uint32_t span = inds[colptr[c + 1] - 1] - inds[colptr[c]];
        if (span > best) best = span;
    }
    return best;
}

/* ------------------------------------------------------------------------- */
/* Helpers shared by tests and the bench baseline.                            */
/* ------------------------------------------------------------------------- */

/* Order-sensitive hash over 8-byte words: h = h*1099511628211 XOR w. */
ORACLE_API uint64_t oracle_hash_words(const uint64_t *w, uint64_t n)
{
    uint64_t h = 0;
    for (uint64_t i = 0; i < n; i++) h = (h * 1099511628211ull) ^ w[i];
    return h;
}

/* HardwareSpMV::compareGolden, software/HardwareSpMV.cpp:37-39: bytewise memcmp. */
ORACLE_API int oracle_compare_golden(const void *golden, const void *y, uint32_t rows)
{
    return memcmp(golden, y, (size_t)rows * 8);
}

/* Per-row bound for the re-associated (merge-path) fp64 sum, SURVEY.md 8(c):      */
/* bound[i] = sum_j |a_ij * x_j| (the caller scales by 2*n_i*2^-53).               */
ORACLE_API void oracle_row_abs_sums_f64(uint32_t rows, uint32_t cols,
                                        const uint32_t *colptr, const uint32_t *rowind,
                                        const double *vals, const double *x,
                                        double *abs_sum, uint32_t *row_nnz)
{
    for (uint32_t r = 0; r < rows; r++) { abs_sum[r] = 0.0; row_nnz[r] = 0; }
    for (uint32_t col = 0; col < cols; col++) {
        double xj = x[col];
        for (uint32_t e = colptr[col]; e < colptr[col + 1]; e++) {
            abs_sum[rowind[e]] += fabs(vals[e] * xj);
            row_nnz[rowind[e]]++;
        }
    }
}

/* Timed CPU baseline: `reps` runs of the reference loop, loop-only bracket like   */
/* SoftwareSpMV.cpp:57,66-67.  Returns best seconds; y is re-zeroed between runs.   */
static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

ORACLE_API double oracle_time_spmv_csc_f64(uint32_t rows, uint32_t cols,
                                           const uint32_t *colptr, const uint32_t *rowind,
                                           const double *vals, const double *x, double *y,
                                           int reps)
{
    double best = 1e300;
    for (int i = 0; i < reps; i++) {
        memset(y, 0, (size_t)rows * 8);
        double t0 = now_s();
        oracle_spmv_csc_f64(rows, cols, colptr, rowind, vals, x, y, 0xFFFFFFFFu);
        double t1 = now_s();
        if (t1 - t0 < best) best = t1 - t0;
    }
    return best;
}

/* Power iteration as configs[4] defines it: y = A*x, x <- y/||y||2. */
/* Built on the reference loop above; returns the last norm.                        */
ORACLE_API double oracle_power_iterate_f64(uint32_t n, const uint32_t *colptr,
                                           const uint32_t *rowind, const double *vals,
                                           double *x, double *y, int steps)
{
    double nrm = 0.0;
    for (int s = 0; s < steps; s++) {
        memset(y, 0, (size_t)n * 8);
        oracle_spmv_csc_f64(n, n, colptr, rowind, vals, x, y, 0xFFFFFFFFu);
        double ss = 0.0;
        for (uint32_t i = 0; i < n; i++) ss += y[i] * y[i];
        nrm = sqrt(ss);
        for (uint32_t i = 0; i < n; i++) x[i] = y[i] / nrm;
    }
    return nrm;
}

/* u64 variant: x <- y, no normalisation (wrapping). */
ORACLE_API void oracle_power_iterate_u64(uint32_t n, const uint32_t *colptr,
                                         const uint32_t *rowind, const uint64_t *vals,
                                         uint64_t *x, uint64_t *y, int steps)
{
    for (int s = 0; s < steps; s++) {
        memset(y, 0, (size_t)n * 8);
        oracle_spmv_csc_u64(n, n, colptr, rowind, vals, x, y, 0xFFFFFFFFu);
        memcpy(x, y, (size_t)n * 8);
    }
}

/* ------------------------------------------------------------------------- */
/* Functional model of the direct-mapped write-back vector cache over y,      */
/* chisel/cache-new/NoWMVectorCache.scala: index/tag split :14-24, hit/miss   */
/* :94-95, fill (all invalid) :197-206, miss handling with eviction of a      */
/* valid line :254-268, cold-miss skip on the row-start bit :240-252, flush   */
/* of the valid lines :208-231.  Sequential, one access after the other.      */
/* out[7]: reads, readMisses, conflictEvictions, coldMissSkips, ddrReads,     */
/* flushWrites, hits.                                                         */
/* PARITY UNPINNED: the reference ships no counter fixtures for this cache    */
/* and its Chisel simulator cannot run here (no Scala toolchain); this is a   */
/* reading of the RTL's state machine, checked against nothing but itself.    */
/* ------------------------------------------------------------------------- */
ORACLE_API void oracle_vector_cache_model(uint32_t nnz, const uint32_t *inds, int depth_log2, int cms, uint64_t *out)
{
    const uint32_t depth = 1u << depth_log2, set_mask = depth - 1u;
    const uint32_t addr_mask = cms ? 0x7FFFFFFFu : 0xFFFFFFFFu;
    uint32_t *tag = (uint32_t *)malloc((size_t)depth * sizeof(uint32_t));
    unsigned char *valid = (unsigned char *)calloc(depth, 1);
    uint64_t miss = 0, conflict = 0, cold = 0, flush = 0;
    for (uint32_t e = 0; e < nnz; e++) {
        const uint32_t a = inds[e] & addr_mask, set = a & set_mask, t = a >> depth_log2;
        if (valid[set] && tag[set] == t) continue;              /* readHit */
        miss++;                                                  /* regReadMissCount */
        if (valid[set]) conflict++;                              /* needEvict: memWrReq */
        if (cms && (inds[e] >> 31)) cold++;                      /* sColdMiss: no DDR read */
        valid[set] = 1; tag[set] = t;                            /* sReadMiss3 */
    }
    for (uint32_t i = 0; i < depth; i++) flush += valid[i];      /* sFlush */
    out[0] = nnz; out[1] = miss; out[2] = conflict; out[3] = cold; out[4] = miss - cold; out[5] = flush; out[6] = nnz - miss;
    free(tag); free(valid);
}
