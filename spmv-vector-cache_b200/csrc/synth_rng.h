// synth_rng.h -- counter-based synthetic matrix definitions shared by the device
// generators (synth.cu) and the host mirror the parity tests use (oracle/synth_host.c).
//
// Every entry of every synthetic matrix is a pure function of (seed, row, k) or
// (seed, edge), so any row can be regenerated on the host without materialising the
// whole matrix (SURVEY.md 7.1 step 5, 8(d) "Synthetic inputs").  The reference has no
// generator; the shapes are the ones BASELINE.json's configs name.
#ifndef B200SPMV_SYNTH_RNG_H_
#define B200SPMV_SYNTH_RNG_H_

#include <stdint.h>

#if defined(__CUDACC__)
#define SYNTH_HD __host__ __device__ __forceinline__
#else
#define SYNTH_HD static inline
#endif

#define SYNTH_K_ROW 0x9E3779B97F4A7C15ull
#define SYNTH_K_IDX 0xC2B2AE3D27D4EB4Full
#define SYNTH_K_VAL 0x165667B19E3779F9ull
#define SYNTH_DEFAULT_SEED 0x5EED5EEDull

// matrix kinds (b200spmv_setup_synthetic `kind`)
#define SYNTH_UNIFORM 0 /* nnz_per_row uniform-random columns per row, sorted, dups kept */
#define SYNTH_RMAT 1    /* Graph500 R-MAT (0.57,0.19,0.19,0.05), edge factor = nnz_per_row */
#define SYNTH_BANDED 2  /* 11-wide band (offsets -5..+5, clipped) + nnz_per_row random cols */
#define SYNTH_BAND 3    /* pure band: nnz_per_row consecutive columns from r - nnz_per_row/2, clipped */

#define SYNTH_BAND_HALF 5

SYNTH_HD uint64_t synth_mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// hash of (seed, a, b): SplitMix64(seed ^ a*K1 ^ b*K2) re-mixed once so that the
// xor of two multiplied counters does not leave lattice structure in the low bits.
SYNTH_HD uint64_t synth_hash(uint64_t seed, uint64_t a, uint64_t b)
{
    return synth_mix64(synth_mix64(seed ^ (a * SYNTH_K_ROW)) ^ (b * SYNTH_K_IDX));
}

SYNTH_HD uint64_t synth_mulhi64(uint64_t a, uint64_t b)
{
#if defined(__CUDA_ARCH__)
    return __umul64hi(a, b);
#else
    return (uint64_t)(((unsigned __int128)a * (unsigned __int128)b) >> 64);
#endif
}

// uniform integer in [0, n)
SYNTH_HD uint32_t synth_below(uint64_t h, uint32_t n) { return (uint32_t)synth_mulhi64(h, (uint64_t)n); }

// uniform double in [0,1) from the top 53 bits
SYNTH_HD double synth_unit(uint64_t h) { return (double)(h >> 11) * (1.0 / 9007199254740992.0); }

// value bits of entry (a, b) for dtype f64 (returned as the IEEE bit pattern) or u64.
//   kind UNIFORM / RMAT : f64 uniform in [-1, 1); u64 = the raw 64 random bits
//   kind BANDED         : band entries 1/16, random entries uniform in [0, 1/16);
//                         u64: band entries 1, random entries raw bits
SYNTH_HD uint64_t synth_value_bits(uint64_t seed, uint64_t a, uint64_t b, int dtype_u64, int kind,
                                   int is_band)
{
    uint64_t h = synth_hash(seed ^ SYNTH_K_VAL, a, b);
    if (dtype_u64) return (kind == SYNTH_BANDED && is_band) ? 1ull : h;
    double v;
    if (kind == SYNTH_BANDED)
        v = is_band ? 0.0625 : synth_unit(h) * 0.0625;
    else
        v = 2.0 * synth_unit(h) - 1.0;
    union { double d; uint64_t u; } cv;
    cv.d = v;
    return cv.u;
}

// x vector entry: uniform [0,1) from seed+1 (f64) or raw bits (u64)
SYNTH_HD uint64_t synth_x_bits(uint64_t seed, uint64_t i, int dtype_u64)
{
    uint64_t h = synth_hash(seed + 1, i, 0x78ull);
    if (dtype_u64) return h;
    union { double d; uint64_t u; } cv;
    cv.d = synth_unit(h);
    return cv.u;
}

// R-MAT edge e -> (row, col) with `scale` levels; 16 random bits per level.
SYNTH_HD void synth_rmat_edge(uint64_t seed, uint64_t e, int scale, uint32_t *row, uint32_t *col)
{
    uint32_t r = 0, c = 0;
    uint64_t word = 0;
    for (int lvl = 0; lvl < scale; lvl++) {
        if ((lvl & 3) == 0) word = synth_hash(seed, e, (uint64_t)(lvl >> 2));
        uint32_t u = (uint32_t)(word & 0xFFFFu);
        word >>= 16;
        // a=0.57 | b=0.19 | c=0.19 | d=0.05  (thresholds * 65536)
        uint32_t rb = (u >= 49807u) ? 1u : 0u;                              // c or d -> lower half
        uint32_t cb = ((u >= 37356u && u < 49807u) || u >= 62259u) ? 1u : 0u; // b or d -> right half
        r = (r << 1) | rb;
        c = (c << 1) | cb;
    }
    *row = r;
    *col = c;
}

// number of band entries of row r in an n x n banded matrix
SYNTH_HD uint32_t synth_band_count(uint32_t r, uint32_t n)
{
    uint32_t lo = r < SYNTH_BAND_HALF ? r : SYNTH_BAND_HALF;
    uint32_t hi = (n - 1 - r) < SYNTH_BAND_HALF ? (n - 1 - r) : SYNTH_BAND_HALF;
    return lo + 1 + hi;
}

#define SYNTH_MAX_NPR 64

// pure band (SYNTH_BAND): row r holds columns [lo, hi] = [r - npr/2, r - npr/2 + npr - 1] clipped to the matrix
SYNTH_HD void synth_pure_band_range(uint32_t r, uint32_t cols, uint32_t npr, uint32_t *lo, uint32_t *hi)
{
    const int64_t first = (int64_t)r - (int64_t)(npr / 2), last = first + (int64_t)npr - 1;
    *lo = first < 0 ? 0u : (uint32_t)first;
    *hi = last >= (int64_t)cols ? cols - 1 : (uint32_t)last;
}

// length of row r for the row-generated kinds
SYNTH_HD uint32_t synth_row_len(int kind, uint32_t r, uint32_t rows, uint32_t npr)
{
    if (kind == SYNTH_BAND) {
        uint32_t lo, hi;
        synth_pure_band_range(r, rows, npr, &lo, &hi);
        return hi - lo + 1;
    }
    return kind == SYNTH_BANDED ? synth_band_count(r, rows) + npr : npr;
}

// Generate row r (kinds UNIFORM and BANDED) in CSR order: columns ascending, duplicates
// kept; for BANDED a band entry precedes a random entry of the same column.
// out_col/out_val need room for synth_row_len() entries.  Returns the row length.
SYNTH_HD uint32_t synth_gen_row(int kind, uint64_t seed, uint32_t r, uint32_t rows, uint32_t cols,
                                uint32_t npr, int dtype_u64, uint32_t *out_col, uint64_t *out_val)
{
    if (kind == SYNTH_BAND) {
        uint32_t lo, hi;
        synth_pure_band_range(r, cols, npr, &lo, &hi);
        for (uint32_t c = lo; c <= hi; c++) {
            out_col[c - lo] = c;
            out_val[c - lo] = synth_value_bits(seed, r, c - lo, dtype_u64, kind, 0);
        }
        return hi - lo + 1;
    }
    uint32_t draws[SYNTH_MAX_NPR];
    for (uint32_t k = 0; k < npr; k++) {
        uint32_t c = synth_below(synth_hash(seed, r, k), cols);
        uint32_t j = k; // insertion sort, ascending
        while (j > 0 && draws[j - 1] > c) {
            draws[j] = draws[j - 1];
            j--;
        }
        draws[j] = c;
    }
    if (kind != SYNTH_BANDED) {
        for (uint32_t p = 0; p < npr; p++) {
            out_col[p] = draws[p];
            out_val[p] = synth_value_bits(seed, r, p, dtype_u64, kind, 0);
        }
        return npr;
    }
    uint32_t lo = r < SYNTH_BAND_HALF ? r : SYNTH_BAND_HALF;
    uint32_t nb = synth_band_count(r, rows);
    uint32_t bi = 0, qi = 0, n = 0;
    while (bi < nb || qi < npr) {
        uint32_t bc = r - lo + bi;
        if (qi >= npr || (bi < nb && bc <= draws[qi])) {
            out_col[n] = bc;
            out_val[n] = synth_value_bits(seed, r, 1000 + bi, dtype_u64, kind, 1);
            bi++;
        } else {
            out_col[n] = draws[qi];
            out_val[n] = synth_value_bits(seed, r, qi, dtype_u64, kind, 0);
            qi++;
        }
        n++;
    }
    return n;
}

#endif
