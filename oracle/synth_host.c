/*
 * synth_host.c -- host mirror of the device synthetic-matrix generators.
 *
 * TEST INFRASTRUCTURE ONLY (same rule as spmv_oracle.c).  The reference ships no
 * generator; BASELINE.json's configs name the shapes.  The definitions live in
 * spmv-vector-cache_b200/csrc/synth_rng.h (pure functions of seed/row/k), so this
 * file regenerates on the CPU, in the reference's CSC layout, exactly the matrix
 * the GPU generated, for the parity tests and the CPU baseline.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "../spmv-vector-cache_b200/csrc/synth_rng.h"

#define ORACLE_API __attribute__((visibility("default")))

/* number of stored entries of the synthetic matrix */
ORACLE_API uint64_t synth_host_nnz(int kind, uint32_t rows, uint32_t npr)
{
    if (kind == SYNTH_BANDED) {
        uint64_t n = 0;
        for (uint32_t r = 0; r < rows; r++) n += synth_band_count(r, rows) + npr;
        return n;
    }
    if (kind == SYNTH_BAND) {
        uint64_t n = 0;
        for (uint32_t r = 0; r < rows; r++) n += synth_row_len(kind, r, rows, npr);
        return n;
    }
    return (uint64_t)rows * npr;
}

/* CSR of the row-generated kinds (UNIFORM, BANDED, BAND).  Rows are independent pure functions of (seed, r),
 * so the row loop is an OpenMP parallel-for when the library is built with -fopenmp (bench.py materialises the
 * 16M x 16M x 32 headline matrix on the host for the reference arm). */
ORACLE_API void synth_host_csr(int kind, uint64_t seed, uint32_t rows, uint32_t cols, uint32_t npr,
                               int dtype_u64, uint32_t *rowptr, uint32_t *colind, uint64_t *vals)
{
    uint32_t p = 0;
    for (uint32_t r = 0; r < rows; r++) {
        rowptr[r] = p;
        p += synth_row_len(kind, r, rows, npr);
    }
    rowptr[rows] = p;
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < (int64_t)rows; r++)
        synth_gen_row(kind, seed, (uint32_t)r, rows, cols, npr, dtype_u64, colind + rowptr[r], vals + rowptr[r]);
}

/* COO edge list of the R-MAT kind, in edge order (rows = cols = 1 << scale). */
ORACLE_API void synth_host_rmat_coo(uint64_t seed, int scale, uint64_t nedges, int dtype_u64,
                                    uint32_t *row, uint32_t *col, uint64_t *vals)
{
    for (uint64_t e = 0; e < nedges; e++) {
        synth_rmat_edge(seed, e, scale, &row[e], &col[e]);
        vals[e] = synth_value_bits(seed, e, 0, dtype_u64, SYNTH_RMAT, 0);
    }
}

ORACLE_API void synth_host_x(uint64_t seed, uint64_t n, int dtype_u64, uint64_t *x)
{
    for (uint64_t i = 0; i < n; i++) x[i] = synth_x_bits(seed, i, dtype_u64);
}

/* Stable counting sort of a COO list by key (key = col gives CSC, key = row gives CSR):
 * the same count / running-sum / cursor-scatter steps as software/csr2csc.c:20-33. */
ORACLE_API void synth_host_coo_to_compressed(uint32_t nkeys, uint64_t nnz, const uint32_t *key,
                                             const uint32_t *other, const uint64_t *vals,
                                             uint32_t *ptr, uint32_t *ind, uint64_t *out_vals)
{
    uint32_t *cur = (uint32_t *)calloc((size_t)nkeys + 1, sizeof(uint32_t));
    for (uint64_t e = 0; e < nnz; e++) cur[key[e] + 1]++;
    for (uint32_t i = 0; i < nkeys; i++) cur[i + 1] += cur[i];
    memcpy(ptr, cur, ((size_t)nkeys + 1) * sizeof(uint32_t));
    for (uint64_t e = 0; e < nnz; e++) {
        uint32_t l = cur[key[e]]++;
        ind[l] = other[e];
        out_vals[l] = vals[e];
    }
    free(cur);
}

/* y[r] of y = A*x for ONE row of a row-generated synthetic matrix with the synthetic x
 * (synth_x_bits(xseed, col)), summed in CSR = reference order with separate mul and add.
 * Lets a test spot-check rows of a matrix far too large to materialise on the host. */
ORACLE_API uint64_t synth_host_row_dot(int kind, uint64_t seed, uint32_t r, uint32_t rows, uint32_t cols,
                                       uint32_t npr, int dtype_u64, uint64_t xseed, double *abs_sum)
{
    uint32_t c[SYNTH_MAX_NPR + 2 * SYNTH_BAND_HALF + 1];
    uint64_t v[SYNTH_MAX_NPR + 2 * SYNTH_BAND_HALF + 1];
    uint32_t n = synth_gen_row(kind, seed, r, rows, cols, npr, dtype_u64, c, v);
    if (dtype_u64) {
        uint64_t acc = 0;
        for (uint32_t k = 0; k < n; k++) acc += v[k] * synth_x_bits(xseed, c[k], 1);
        return acc;
    }
    union { double d; uint64_t u; } a, b, out;
    double acc = 0.0, mag = 0.0;
    for (uint32_t k = 0; k < n; k++) {
        a.u = v[k];
        b.u = synth_x_bits(xseed, c[k], 0);
        double prod = a.d * b.d;
        acc += prod;
        mag += prod < 0 ? -prod : prod;
    }
    if (abs_sum) *abs_sum = mag;
    out.d = acc;
    return out.u;
}
