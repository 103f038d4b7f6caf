// spmv.cu -- the hot path: y (+)= A*x on the device-resident CSR (software/SoftwareSpMV.cpp:59-64
// after the CSC->CSR transpose).
//
//   spmv_merge_kernel : nnz-balanced merge-path SpMV.  The merge of the row-end offsets with
//       the nnz indices is cut into tiles of MP_TILE items; a CTA stages one tile's products
//       val*x[col] in shared memory with 128-bit streaming loads of indices and values, then
//       every thread walks MP_IPT merge items, rows that straddle threads are completed from
//       the neighbours' carries, rows that straddle tiles through a carry-out fixed up by
//       spmv_fixup_kernel.  Re-associates fp64 sums (tolerance mode); u64 is exact.
//   spmv_rowseq_kernel : one thread per row, entries in CSR (= reference) order, separate
//       multiply and add: fp64 results memcmp-equal to SoftwareSpMV (strict mode).
//
// Column slabs: with S > 1 the CSR is stored slab-major and the kernels run slab by slab, so
// the x slice gathered at any time fits the L2 set-aside (the "vector cache").
#include <algorithm>
#include <vector>

#include "common.cuh"

namespace {

constexpr int MP_THREADS = 256;
constexpr int MP_IPT = 8;
constexpr int MP_TILE = MP_THREADS * MP_IPT;  // merge items per tile
constexpr uint32_t NO_ROW = 0xFFFFFFFFu;

// ---- merge-path partition (once per matrix) --------------------------------------------
// rowptr: this slab's [rows+1] absolute offsets.  Tile t of the slab starts at merge diagonal
// t*MP_TILE; writes (rows consumed, absolute nz offset) for t in [0, ntiles].
__global__ void tile_search_kernel(const uint32_t *__restrict__ rowptr, uint32_t rows, uint32_t ntiles,
                                   uint32_t *__restrict__ tile_row, uint32_t *__restrict__ tile_nz)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > ntiles) return;
    const uint32_t nz_begin = rowptr[0];
    const uint64_t nnz = (uint64_t)rowptr[rows] - nz_begin;
    const uint64_t items = (uint64_t)rows + nnz;
    const uint64_t diag = min((unsigned long long)t * MP_TILE, (unsigned long long)items);
    uint64_t lo = diag > nnz ? diag - nnz : 0;
    uint64_t hi = min((unsigned long long)diag, (unsigned long long)rows);
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        // row-end item `mid` precedes nnz item (diag-mid-1) iff rowend[mid] <= that nnz index
        if ((uint64_t)(rowptr[mid + 1] - nz_begin) <= diag - mid - 1) lo = mid + 1; else hi = mid;
    }
    tile_row[t] = (uint32_t)lo;
    tile_nz[t] = nz_begin + (uint32_t)(diag - lo);
}

template <bool PERSIST>
__device__ __forceinline__ uint64_t load_x(const uint64_t *__restrict__ x, uint32_t col, uint64_t pol)
{
    if (PERSIST) return ld_gather_u64(x + col, pol);
    return __ldg(reinterpret_cast<const unsigned long long *>(x + col));
}

struct MergeSmem {
    uint64_t prod[MP_TILE + 4];
    uint64_t yrow[MP_TILE];
    uint64_t cval[MP_THREADS];
    uint32_t rowend[MP_TILE + 1];
    uint32_t crow[MP_THREADS];
};

// ---- merge-path SpMV, one tile per CTA ---------------------------------------------------
// READ_Y: the rows completed here start from the y already in memory (accumulate mode, or any
// slab after the first); otherwise from the additive identity.
template <int DT, bool PERSIST, bool READ_Y>
__global__ void __launch_bounds__(MP_THREADS)
spmv_merge_kernel(const uint32_t *__restrict__ rowptr, const uint32_t *__restrict__ colind,
                  const uint64_t *__restrict__ vals, const uint64_t *__restrict__ x,
                  uint64_t *__restrict__ y, uint32_t rows, const uint32_t *__restrict__ tile_row,
                  const uint32_t *__restrict__ tile_nz, uint32_t *__restrict__ carry_row,
                  uint64_t *__restrict__ carry_val, uint32_t colmask)
{
    __shared__ __align__(16) MergeSmem sm;
    const uint32_t t = blockIdx.x;
    const int tid = threadIdx.x;
    const uint32_t r0 = tile_row[t], r1 = tile_row[t + 1];
    const uint32_t nz0 = tile_nz[t], nz1 = tile_nz[t + 1];
    const uint32_t tile_rows = r1 - r0, tile_nnz = nz1 - nz0;
    const uint32_t tile_items = tile_rows + tile_nnz;

    // (1) row-end offsets of rows r0..r1 (r1 = the row left open at the end of the tile)
    for (uint32_t i = tid; i <= tile_rows; i += MP_THREADS) {
        const uint32_t r = r0 + i;
        sm.rowend[i] = (r < rows) ? rowptr[r + 1] : 0xFFFFFFFFu;
    }
    // (2) stage products: groups of 4 entries from a 16-byte aligned base
    const uint32_t base = nz0 & ~3u;
    const uint32_t ngroups = (nz1 - base + 3) >> 2;
    const uint64_t pol_stream = policy_evict_first(), pol_keep = policy_evict_last();
    for (uint32_t g = tid; g < ngroups; g += MP_THREADS) {
        const uint64_t e = (uint64_t)base + 4ull * g;
        const uint4 c = ld_stream_v4(colind + e, pol_stream);
        uint64_t vv[4];
        ld_stream_v4_b64(vals + e, vv);
        const uint32_t cc[4] = {c.x & colmask, c.y & colmask, c.z & colmask, c.w & colmask};
        uint64_t xv[4];
#pragma unroll
        for (int k = 0; k < 4; k++) xv[k] = (e + k < nz1) ? load_x<PERSIST>(x, cc[k], pol_keep) : 0ull;
#pragma unroll
        for (int k = 0; k < 4; k++) sm.prod[4 * g + k] = Semi<DT>::mul(vv[k], xv[k]);
    }
    __syncthreads();

    // (3) each thread walks MP_IPT merge items starting at its diagonal
    const uint32_t diag = min((uint32_t)tid * MP_IPT, tile_items);
    uint32_t lo = diag > tile_nnz ? diag - tile_nnz : 0u;
    uint32_t hi = min(diag, tile_rows);
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (sm.rowend[mid] <= nz0 + (diag - mid - 1)) lo = mid + 1; else hi = mid;
    }
    uint32_t r = lo;                    // local row
    uint32_t nz = nz0 + (diag - lo);    // absolute nz
    const uint32_t first_row = r;
    bool emitted = false;
    uint64_t acc = 0;
    uint32_t row_end = sm.rowend[r];
#pragma unroll
    for (int k = 0; k < MP_IPT; k++) {
        if (diag + k < tile_items) {
            if (nz < row_end) {
                acc = Semi<DT>::add(acc, sm.prod[nz - base]);
                nz++;
            } else {
                sm.yrow[r] = acc;
                emitted = true;
                acc = 0;
                r++;
                row_end = sm.rowend[r];
            }
        }
    }
    sm.crow[tid] = r;
    sm.cval[tid] = acc;
    __syncthreads();

    // (4) the first row a thread completed may have begun in earlier threads: add their carries
    if (emitted) {
        int k = tid;
        while (k > 0 && sm.crow[k - 1] == first_row) k--;
        if (k < tid) {
            uint64_t c = sm.cval[k];
            for (int q = k + 1; q < tid; q++) c = Semi<DT>::add(c, sm.cval[q]);
            sm.yrow[first_row] = Semi<DT>::add(c, sm.yrow[first_row]);
        }
    }
    // tile carry-out: the partial of the row left open (threads whose walk ended in it)
    if (tid == MP_THREADS - 1) {
        const bool has_items = tile_rows > 0 ? (nz1 > sm.rowend[tile_rows - 1]) : (tile_nnz > 0);
        int k = tid;
        while (k > 0 && sm.crow[k - 1] == tile_rows) k--;
        uint64_t c = sm.cval[k];
        for (int q = k + 1; q <= tid; q++) c = Semi<DT>::add(c, sm.cval[q]);
        carry_row[t] = has_items ? r1 : NO_ROW;
        carry_val[t] = c;
    }
    __syncthreads();

    // (5) coalesced write of the rows completed in this tile
    for (uint32_t i = tid; i < tile_rows; i += MP_THREADS) {
        uint64_t v = sm.yrow[i];
        if (READ_Y) v = Semi<DT>::add(y[r0 + i], v);
        y[r0 + i] = v;
    }
}

// ---- merge-path SpMV v2: TMA-staged tile, register accumulation ----------------------------
// One CTA per tile, like v1, but (a) the tile's column indices, values and row-end offsets are
// brought into shared memory by three 1-D bulk async copies (TMA, cp.async.bulk) issued by one
// thread and signalled on an mbarrier -- no LSU wavefronts and no registers are spent on the
// streams, and with 6 CTAs resident per SM ~200 KB of stream data is in flight per SM;
// (b) every thread resolves its MP_IPT merge items first (which are entries, which are row
// ends), issues all its x gathers back to back, then accumulates in registers: products never
// go through shared memory; (c) completed rows are stored straight to y.
// The LSU is left to what bounds this kernel on a matrix without locality: one L2 sector
// request per gathered x element (profiles/r01_gather_ceiling.txt).
struct TmaSmem {
    uint64_t val[MP_TILE + 8];
    uint64_t cval[MP_THREADS];
    uint32_t col[MP_TILE + 8];
    uint32_t rowend[MP_TILE + 12];
    uint32_t crow[MP_THREADS];
    uint64_t mbar;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tma_load_1d(void *dst_smem, const void *src, uint32_t bytes, uint64_t *mbar, uint64_t pol)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
        ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(mbar)), "l"(pol) : "memory");
}

template <int DT, bool PERSIST, bool READ_Y>
__global__ void __launch_bounds__(MP_THREADS, 6)
spmv_merge_tma_kernel(const uint32_t *__restrict__ rowptr, const uint32_t *__restrict__ colind,
                      const uint64_t *__restrict__ vals, const uint64_t *__restrict__ x,
                      uint64_t *__restrict__ y, uint32_t rows, const uint32_t *__restrict__ tile_row,
                      const uint32_t *__restrict__ tile_nz, uint32_t *__restrict__ carry_row,
                      uint64_t *__restrict__ carry_val, uint32_t colmask)
{
    __shared__ __align__(128) TmaSmem sm;
    const uint32_t t = blockIdx.x;
    const int tid = threadIdx.x;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&sm.mbar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const uint32_t r0 = __ldg(tile_row + t), r1 = __ldg(tile_row + t + 1);
    const uint32_t nz0 = __ldg(tile_nz + t), nz1 = __ldg(tile_nz + t + 1);
    const uint32_t tile_rows = r1 - r0, tile_nnz = nz1 - nz0;
    const uint32_t tile_items = tile_rows + tile_nnz;
    const uint32_t base = nz0 & ~3u;                       // 16-byte aligned start of the entry streams
    const uint32_t cnt = (nz1 - base + 3u) & ~3u;          // entries copied (multiple of 4)
    const uint32_t rbase = (r0 + 1u) & ~3u;                // rowptr[r0+1 .. r1+1] = row ends of rows r0..r1
    const uint32_t rcnt = (r1 + 2u - rbase + 3u) & ~3u;
    const uint32_t roff = r0 + 1u - rbase;
    __syncthreads();
    if (tid == 0) {
        const uint64_t pol = policy_evict_first();
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&sm.mbar)),
                     "r"(cnt * 12u + rcnt * 4u) : "memory");
        tma_load_1d(sm.rowend, rowptr + rbase, rcnt * 4u, &sm.mbar, pol);
        if (cnt) {
            tma_load_1d(sm.col, colind + base, cnt * 4u, &sm.mbar, pol);
            tma_load_1d(sm.val, vals + base, cnt * 8u, &sm.mbar, pol);
        }
    }
    {   // wait for the three copies (phase 0)
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(smem_u32(&sm.mbar)) : "memory");
    }
    const uint32_t *rowend = sm.rowend + roff;             // rowend[i] = end offset of local row i

    // merge-path start of this thread inside the tile
    const uint32_t diag = min((uint32_t)tid * MP_IPT, tile_items);
    uint32_t lo = diag > tile_nnz ? diag - tile_nnz : 0u;
    uint32_t hi = min(diag, tile_rows);
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (rowend[mid] <= nz0 + (diag - mid - 1)) lo = mid + 1; else hi = mid;
    }
    const uint32_t first_row = lo;
    const uint32_t nz_first = nz0 + (diag - lo);
    // pass 1: which of my items are row ends (bit k), how many are entries
    uint32_t r = lo, nz = nz_first, ends = 0;
    uint32_t row_end = rowend[r];
#pragma unroll
    for (int k = 0; k < MP_IPT; k++) {
        if (diag + k < tile_items) {
            if (nz < row_end) nz++;
            else { ends |= 1u << k; r++; row_end = rowend[r]; }
        }
    }
    // pass 2: all gathers in flight together.  Item k, when it is an entry, is my
    // (k - #row-ends before k)-th entry.
    const uint64_t pol_keep = policy_evict_last();
    const uint32_t off = nz_first - base;
    uint64_t xv[MP_IPT];
#pragma unroll
    for (int k = 0; k < MP_IPT; k++) {
        const bool is_entry = (diag + k < tile_items) && !((ends >> k) & 1u);
        const uint32_t jk = k - __popc(ends & ((1u << k) - 1u));
        xv[k] = is_entry ? load_x<PERSIST>(x, sm.col[off + jk] & colmask, pol_keep) : 0ull;
    }
    // pass 3: accumulate in merge order
    uint64_t acc = 0, first_val = 0;
    bool emitted = false;
    uint32_t rr = first_row;
#pragma unroll
    for (int k = 0; k < MP_IPT; k++) {
        if (diag + k < tile_items) {
            if ((ends >> k) & 1u) {
                if (!emitted) { first_val = acc; emitted = true; }
                else {
                    uint64_t v = acc;
                    if (READ_Y) v = Semi<DT>::add(y[r0 + rr], v);
                    y[r0 + rr] = v;
                }
                acc = 0;
                rr++;
            } else {
                const uint32_t jk = k - __popc(ends & ((1u << k) - 1u));
                acc = Semi<DT>::add(acc, Semi<DT>::mul(sm.val[off + jk], xv[k]));
            }
        }
    }
    sm.crow[tid] = rr;
    sm.cval[tid] = acc;
    __syncthreads();

    if (emitted) {  // my first completed row may have begun in earlier threads
        int k = tid;
        while (k > 0 && sm.crow[k - 1] == first_row) k--;
        uint64_t v = first_val;
        if (k < tid) {
            uint64_t c = sm.cval[k];
            for (int q = k + 1; q < tid; q++) c = Semi<DT>::add(c, sm.cval[q]);
            v = Semi<DT>::add(c, v);
        }
        if (READ_Y) v = Semi<DT>::add(y[r0 + first_row], v);
        y[r0 + first_row] = v;
    }
    if (tid == MP_THREADS - 1) {  // tile carry-out: partial of the row left open
        const bool has_items = tile_rows > 0 ? (nz1 > rowend[tile_rows - 1]) : (tile_nnz > 0);
        int k = tid;
        while (k > 0 && sm.crow[k - 1] == tile_rows) k--;
        uint64_t c = sm.cval[k];
        for (int q = k + 1; q <= tid; q++) c = Semi<DT>::add(c, sm.cval[q]);
        carry_row[t] = has_items ? r1 : NO_ROW;
        carry_val[t] = c;
    }
}

// ---- kernel 3: flagged segmented sum -------------------------------------------------------
// The CSR built at setup carries a row-start flag in bit 31 of every row's first column index,
// and each slab's entries are padded to whole tiles (the first pad entry is flagged: it closes
// the last row).  Stored entries are then split EXACTLY evenly: every thread owns 8 consecutive
// entries, loaded with 256-bit loads straight into registers (one 32-byte sector of indices,
// two of values), gathers its 8 x elements, and sums them sequentially, cutting at the flags.
// Threads are stitched together by a warp-shuffle segmented scan of their open tails, warps
// through shared memory, tiles through the carry arrays + spmv_fixup_kernel.  No row pointers are
// read: seg_row[] (only when a slab has empty rows) maps the n-th flag to its row.
// This is the merge-path idea (nnz-balanced, rows cut wherever the split falls) with the row
// search replaced by flags precomputed at setup; it costs ~6x fewer instructions per entry than
// the searching kernels, which matters because the x gathers leave little issue/LSU headroom.
constexpr int FK_THREADS = 256;
constexpr int FK_WARPS = FK_THREADS / 32;

// 8 consecutive u32 / u64 with 256-bit loads.  L1 allocation stays on: a 256-bit load reaches L2
// as two 128-bit requests for the same 32-byte sector and only an L1 line lets the second merge.
__device__ __forceinline__ void ld256_u32(const void *p, uint32_t *c)
{
    asm volatile("ld.global.nc.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]), "=r"(c[4]), "=r"(c[5]), "=r"(c[6]), "=r"(c[7])
                 : "l"(p));
}
__device__ __forceinline__ void ld256_u64(const void *p, uint64_t *v)
{
    asm volatile("ld.global.nc.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3])
                 : "l"(p));
}

template <int DT, bool READ_Y, bool HAS_EMPTY>
__device__ __forceinline__ void store_segment(uint64_t *__restrict__ y, const uint32_t *__restrict__ seg_row,
                                              uint32_t rows, uint32_t seg, uint64_t v)
{
    const uint32_t row = HAS_EMPTY ? __ldg(seg_row + seg) : seg;
    if (row < rows) {
        if (READ_Y) v = Semi<DT>::add(y[row], v);
        y[row] = v;
    }
}

// EPT = consecutive entries per thread (8 or 16); the tile is always B2_FLAG_TILE entries, so the
// CTA has B2_FLAG_TILE / EPT threads.
template <int DT, bool PERSIST, bool READ_Y, bool HAS_EMPTY, int EPT>
__global__ void __launch_bounds__(B2_FLAG_TILE / EPT, EPT == 8 ? 4 : 4)
spmv_flag_kernel(const uint32_t *__restrict__ colind, const uint64_t *__restrict__ vals,
                 const uint64_t *__restrict__ x, uint64_t *__restrict__ y,
                 const uint32_t *__restrict__ seg_row, uint32_t rows, const uint32_t *__restrict__ tile_seg,
                 uint32_t *__restrict__ carry_row, uint64_t *__restrict__ carry_val, uint32_t pf_dist)
{
    constexpr int THREADS = B2_FLAG_TILE / EPT, WARPS = THREADS / 32;
    // one thread asks the TMA unit to pull the streams of the tile `pf_dist` CTAs ahead into L2 (two bulk
    // prefetches, no registers, no data returned to the SM): that CTA then starts on L2 hits instead of HBM
    if (pf_dist != 0 && threadIdx.x == 0) {
        const uint64_t tp = (uint64_t)blockIdx.x + pf_dist;
        if (tp < gridDim.x) {
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(vals + tp * B2_FLAG_TILE), "r"(B2_FLAG_TILE * 8u) : "memory");
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(colind + tp * B2_FLAG_TILE), "r"(B2_FLAG_TILE * 4u) : "memory");
        }
    }
    __shared__ __align__(16) uint32_t s_wcnt[WARPS];
    __shared__ uint32_t s_wflag[WARPS];
    __shared__ uint64_t s_wsum[WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t t = blockIdx.x;
    const uint64_t pos = (uint64_t)t * B2_FLAG_TILE + (uint64_t)tid * EPT;  // colind/vals are slab-relative

    uint32_t c[EPT];
    uint64_t p[EPT];  // values, then products
#pragma unroll
    for (int g = 0; g < EPT / 8; g++) ld256_u32(colind + pos + 8 * g, c + 8 * g);
#pragma unroll
    for (int g = 0; g < EPT / 4; g++) ld256_u64(vals + pos + 4 * g, p + 4 * g);
    const uint64_t pol_keep = policy_evict_last();
    uint64_t xv[EPT];
#pragma unroll
    for (int i = 0; i < EPT; i++) xv[i] = load_x<PERSIST>(x, c[i] & 0x7FFFFFFFu, pol_keep);

    // row-start flags of my entries
    uint32_t f = 0;
#pragma unroll
    for (int i = 0; i < EPT; i++) f |= (c[i] >> 31) << i;
    const uint32_t nfl = __popc(f);
    const uint32_t lt = (1u << lane) - 1u;
    // flags in lower lanes: sum over the bits of nfl of ballots (no dependent shuffle chain)
    uint32_t lane_prefix = 0, warp_flags = 0;
#pragma unroll
    for (int bit = 0; (1 << bit) <= EPT; bit++) {
        const uint32_t bal_b = __ballot_sync(0xffffffffu, (nfl >> bit) & 1u);
        lane_prefix += __popc(bal_b & lt) << bit;
        warp_flags += __popc(bal_b) << bit;
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, nfl != 0);
    if (lane == 31) s_wcnt[warp] = warp_flags;
    __syncthreads();  // flag counts only: cheap, and everything below overlaps the gathers in flight
    uint32_t segs_before = __ldg(tile_seg + t) + lane_prefix, tile_flags = 0;
    static_assert(WARPS % 4 == 0, "warp counts are read four at a time");
#pragma unroll
    for (int q = 0; q < WARPS / 4; q++) {  // one 128-bit shared load per four warps instead of a load per warp
        const uint4 c4 = reinterpret_cast<const uint4 *>(s_wcnt)[q];
        const uint32_t c[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (q * 4 + j < warp) segs_before += c[j];
            tile_flags += c[j];
        }
    }
    // the row my first flag closes, and its current y (accumulate mode): fetched now, needed last
    uint32_t close_row = NO_ROW;
    uint64_t y_old = 0;
    if (nfl != 0 && segs_before != 0) {
        close_row = HAS_EMPTY ? __ldg(seg_row + segs_before - 1) : segs_before - 1;
        if (close_row >= rows) close_row = NO_ROW;
        if (READ_Y && close_row != NO_ROW) y_old = y[close_row];
    }
    // a second flag in my entries closes a row that lies wholly inside them: its row id and y are fetched
    // now as well (rows shorter than 8 entries are common once a matrix is cut into column slabs)
    uint32_t row2 = NO_ROW;
    uint64_t y_old2 = 0;
    if (nfl > 1) {
        row2 = HAS_EMPTY ? __ldg(seg_row + segs_before) : segs_before;
        if (row2 >= rows) row2 = NO_ROW;
        if (READ_Y && row2 != NO_ROW) y_old2 = y[row2];
    }

    // branch-free sequential pass: head = sum before my first flag, acc = sum after my last flag
    uint64_t acc = 0, head = 0, mid = 0;  // mid = the row between my first and second flag
    uint32_t seen = 0;
#pragma unroll
    for (int i = 0; i < EPT; i++) {
        p[i] = Semi<DT>::mul(p[i], xv[i]);
        const uint32_t fl = (f >> i) & 1u;
        head = (fl && seen == 0) ? acc : head;
        mid = (fl && seen == 1) ? acc : mid;
        seen += fl;
        acc = fl ? 0ull : acc;
        acc = Semi<DT>::add(acc, p[i]);
    }
    // segmented inclusive scan of the open tails over the warp: S = sum of tails from the nearest
    // flagged lane at or below me (or lane 0) to me
    const uint32_t below = bal & ((2u << lane) - 1u);
    const int start = below ? 31 - __clz(below) : 0;
    uint64_t S = acc;
    // steps whose distance exceeds the longest open run of the warp change nothing: skip them (one REDUX
    // instead of up to 8 shuffles when the rows are short, which column slabs make them)
    const int span = __reduce_max_sync(0xffffffffu, lane - start);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        if (d > span) break;
        const uint64_t n = __shfl_up_sync(0xffffffffu, S, d);
        if (lane - d >= start) S = Semi<DT>::add(n, S);
    }
    uint64_t carry_in = __shfl_up_sync(0xffffffffu, S, 1);  // open partial ending at lane-1
    if (lane == 0) carry_in = 0;
    const bool open_left = (bal & lt) == 0;  // no flagged lane below me in this warp
    if (lane == 31) { s_wsum[warp] = S; s_wflag[warp] = bal != 0; }
    __syncthreads();

    // partial carried into this warp: sums of the preceding warps back to the nearest flagged one
    uint64_t cw = 0;
    if (warp > 0) {
        int w0 = warp - 1;
        while (w0 > 0 && !s_wflag[w0]) w0--;
        cw = s_wsum[w0];
        for (int w = w0 + 1; w < warp; w++) cw = Semi<DT>::add(cw, s_wsum[w]);
    }
    if (nfl != 0) {
        if (close_row != NO_ROW) {  // my first flag closes the row that was open on entry
            uint64_t val = open_left ? Semi<DT>::add(cw, carry_in) : carry_in;
            val = Semi<DT>::add(val, head);
            if (READ_Y) val = Semi<DT>::add(y_old, val);
            y[close_row] = val;
        }
        if (nfl > 1 && row2 != NO_ROW) y[row2] = READ_Y ? Semi<DT>::add(y_old2, mid) : mid;
        if (nfl > 2) {  // further rows that begin and end inside my entries (very short rows): re-walk
            uint64_t a = 0;
            uint32_t k = 0;
#pragma unroll
            for (int i = 0; i < EPT; i++) {
                if ((f >> i) & 1u) {
                    if (k > 1) store_segment<DT, READ_Y, HAS_EMPTY>(y, seg_row, rows, segs_before + k - 1, a);
                    k++;
                    a = 0;
                }
                a = Semi<DT>::add(a, p[i]);
            }
        }
    }
    if (tid == THREADS - 1) {  // tile carry-out: the row left open at the end of the tile
        const uint64_t out = bal ? S : Semi<DT>::add(cw, S);
        const uint32_t seg_total = __ldg(tile_seg + t) + tile_flags;
        uint32_t row = NO_ROW;
        if (seg_total != 0) {
            row = HAS_EMPTY ? __ldg(seg_row + seg_total - 1) : seg_total - 1;
            if (row >= rows) row = NO_ROW;
        }
        carry_row[t] = row;
        carry_val[t] = out;
    }
}

// ---- kernel 4: persistent flagged segmented sum with a TMA-fed shared-memory ring ------------------
// Same arithmetic and the same flagged tiles as kernel 3, but the CTAs are persistent (grid = SMs x
// resident CTAs, tiles dealt round-robin) and the two entry streams of tile i + RING are brought into
// a shared-memory ring by 1-D bulk async copies (TMA, cp.async.bulk -> UBLKCP) issued by one thread
// as soon as the ring slot has been read into registers.  What bounds this kernel on a matrix without
// locality is the rate at which one SM can post L2 requests for the x gathers (ncu:
// l1tex__m_l1tex2xbar_req_cycles_active, profiles/r01_k3_flag_ncu_summary.txt); kernel 3 leaves that
// port idle while a fresh CTA waits for its first HBM loads, here the streams are already on chip
// when a tile starts, so the gathers are issued back to back.
constexpr int PK_RING = 2;
constexpr uint32_t PK_STAGE_BYTES = B2_FLAG_TILE * 12u;
constexpr uint32_t PK_SMEM_BYTES = PK_RING * PK_STAGE_BYTES + 128;  // + full[RING], empty[RING] mbarriers

template <int DT, bool PERSIST, bool READ_Y, bool HAS_EMPTY>
__global__ void __launch_bounds__(B2_FLAG_TILE / 8, 4)
spmv_flag_ring_kernel(const uint32_t *__restrict__ colind, const uint64_t *__restrict__ vals,
                      const uint64_t *__restrict__ x, uint64_t *__restrict__ y,
                      const uint32_t *__restrict__ seg_row, uint32_t rows, const uint32_t *__restrict__ tile_seg,
                      uint32_t *__restrict__ carry_row, uint64_t *__restrict__ carry_val, uint32_t ntiles)
{
    constexpr int EPT = 8, THREADS = B2_FLAG_TILE / EPT, WARPS = THREADS / 32;
    extern __shared__ __align__(128) unsigned char pk_smem[];
    // slot s: values at s*PK_STAGE_BYTES (16 KB), column indices behind them (8 KB)
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(pk_smem + PK_RING * PK_STAGE_BYTES);
    __shared__ uint32_t s_wcnt[WARPS];
    __shared__ uint32_t s_wflag[WARPS];
    __shared__ uint64_t s_wsum[WARPS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t pol_stream = policy_evict_first(), pol_keep = policy_evict_last();

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < PK_RING; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar[s])));            // full: producer + tx bytes
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&s_bar[PK_RING + s])), "r"(WARPS));  // empty: one arrival per warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
        for (int s = 0; s < PK_RING; s++) {
            const uint64_t t = (uint64_t)blockIdx.x + (uint64_t)s * gridDim.x;
            if (t < ntiles) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&s_bar[s])), "r"(PK_STAGE_BYTES) : "memory");
                tma_load_1d(pk_smem + s * PK_STAGE_BYTES, vals + t * B2_FLAG_TILE, B2_FLAG_TILE * 8u, &s_bar[s], pol_stream);
                tma_load_1d(pk_smem + s * PK_STAGE_BYTES + B2_FLAG_TILE * 8u, colind + t * B2_FLAG_TILE, B2_FLAG_TILE * 4u, &s_bar[s], pol_stream);
            }
        }
    }
    __syncthreads();

    uint32_t it = 0;
    for (uint64_t t64 = blockIdx.x; t64 < ntiles; t64 += gridDim.x, it++) {
        const uint32_t t = (uint32_t)t64;
        const int slot = it % PK_RING;
        const uint32_t parity = (it / PK_RING) & 1u;
        const uint32_t tseg = __ldg(tile_seg + t);
        {
            uint32_t done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(smem_u32(&s_bar[slot])), "r"(parity) : "memory");
        }
        uint32_t c[EPT];
        uint64_t p[EPT];
        {
            const uint4 *sc = reinterpret_cast<const uint4 *>(pk_smem + slot * PK_STAGE_BYTES + B2_FLAG_TILE * 8u) + tid * 2;
            const uint4 c0 = sc[0], c1 = sc[1];
            c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
        }
        uint64_t xv[EPT];
#pragma unroll
        for (int i = 0; i < EPT; i++) xv[i] = load_x<PERSIST>(x, c[i] & 0x7FFFFFFFu, pol_keep);
        {
            const ulonglong2 *sv = reinterpret_cast<const ulonglong2 *>(pk_smem + slot * PK_STAGE_BYTES) + tid * 4;
#pragma unroll
            for (int g = 0; g < 4; g++) { const ulonglong2 v = sv[g]; p[2 * g] = v.x; p[2 * g + 1] = v.y; }
        }
        // release the slot: my reads (generic proxy) are ordered before the refill (async proxy) that
        // thread 0 issues once every warp has arrived.  (A __syncthreads is not enough here: BAR.SYNC
        // defers its blocking to the next instruction that touches shared memory, and the bulk copy
        // is not one.)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&s_bar[PK_RING + slot])) : "memory");

        uint32_t f = 0;
#pragma unroll
        for (int i = 0; i < EPT; i++) f |= (c[i] >> 31) << i;
        const uint32_t nfl = __popc(f);
        const uint32_t lt = (1u << lane) - 1u;
        uint32_t lane_prefix = 0, warp_flags = 0;
#pragma unroll
        for (int bit = 0; (1 << bit) <= EPT; bit++) {
            const uint32_t bal_b = __ballot_sync(0xffffffffu, (nfl >> bit) & 1u);
            lane_prefix += __popc(bal_b & lt) << bit;
            warp_flags += __popc(bal_b) << bit;
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, nfl != 0);
        if (lane == 31) s_wcnt[warp] = warp_flags;
        __syncthreads();  // (A)
        if (tid == 0) {
            const uint64_t tn = t64 + (uint64_t)PK_RING * gridDim.x;
            if (tn < ntiles) {
                uint32_t done = 0;  // every warp holds its part of the slot in registers
                while (!done)
                    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                                 : "=r"(done) : "r"(smem_u32(&s_bar[PK_RING + slot])), "r"(parity) : "memory");
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&s_bar[slot])), "r"(PK_STAGE_BYTES) : "memory");
                tma_load_1d(pk_smem + slot * PK_STAGE_BYTES, vals + tn * B2_FLAG_TILE, B2_FLAG_TILE * 8u, &s_bar[slot], pol_stream);
                tma_load_1d(pk_smem + slot * PK_STAGE_BYTES + B2_FLAG_TILE * 8u, colind + tn * B2_FLAG_TILE, B2_FLAG_TILE * 4u, &s_bar[slot], pol_stream);
            }
        }
        uint32_t segs_before = tseg + lane_prefix, tile_flags = 0;
#pragma unroll
        for (int w = 0; w < WARPS; w++) {
            const uint32_t n = s_wcnt[w];
            if (w < warp) segs_before += n;
            tile_flags += n;
        }
        uint32_t close_row = NO_ROW;
        uint64_t y_old = 0;
        if (nfl != 0 && segs_before != 0) {
            close_row = HAS_EMPTY ? __ldg(seg_row + segs_before - 1) : segs_before - 1;
            if (close_row >= rows) close_row = NO_ROW;
            if (READ_Y && close_row != NO_ROW) y_old = y[close_row];
        }

        uint64_t acc = 0, head = 0;
        uint32_t seen = 0;
#pragma unroll
        for (int i = 0; i < EPT; i++) {
            p[i] = Semi<DT>::mul(p[i], xv[i]);
            const uint32_t fl = (f >> i) & 1u;
            head = (fl & ~seen & 1u) ? acc : head;
            seen |= fl;
            acc = fl ? 0ull : acc;
            acc = Semi<DT>::add(acc, p[i]);
        }
        const uint32_t below = bal & ((2u << lane) - 1u);
        const int start = below ? 31 - __clz(below) : 0;
        uint64_t S = acc;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint64_t n = __shfl_up_sync(0xffffffffu, S, d);
            if (lane - d >= start) S = Semi<DT>::add(n, S);
        }
        uint64_t carry_in = __shfl_up_sync(0xffffffffu, S, 1);
        if (lane == 0) carry_in = 0;
        const bool open_left = (bal & lt) == 0;
        if (lane == 31) { s_wsum[warp] = S; s_wflag[warp] = bal != 0; }
        __syncthreads();  // (B)

        uint64_t cw = 0;
        if (warp > 0) {
            int w0 = warp - 1;
            while (w0 > 0 && !s_wflag[w0]) w0--;
            cw = s_wsum[w0];
            for (int w = w0 + 1; w < warp; w++) cw = Semi<DT>::add(cw, s_wsum[w]);
        }
        if (nfl != 0) {
            if (close_row != NO_ROW) {
                uint64_t val = open_left ? Semi<DT>::add(cw, carry_in) : carry_in;
                val = Semi<DT>::add(val, head);
                if (READ_Y) val = Semi<DT>::add(y_old, val);
                y[close_row] = val;
            }
            if (nfl > 1) {
                uint64_t a = 0;
                uint32_t k = 0;
#pragma unroll
                for (int i = 0; i < EPT; i++) {
                    if ((f >> i) & 1u) {
                        if (k != 0) store_segment<DT, READ_Y, HAS_EMPTY>(y, seg_row, rows, segs_before + k - 1, a);
                        k++;
                        a = 0;
                    }
                    a = Semi<DT>::add(a, p[i]);
                }
            }
        }
        if (tid == THREADS - 1) {
            const uint64_t out = bal ? S : Semi<DT>::add(cw, S);
            const uint32_t seg_total = tseg + tile_flags;
            uint32_t row = NO_ROW;
            if (seg_total != 0) {
                row = HAS_EMPTY ? __ldg(seg_row + seg_total - 1) : seg_total - 1;
                if (row >= rows) row = NO_ROW;
            }
            carry_row[t] = row;
            carry_val[t] = out;
        }
        // s_wcnt is rewritten before (A) of the next tile and was last read before (B) of this one;
        // s_wsum / s_wflag are rewritten after (A) of the next tile, which every reader above has passed
    }
}

// ---- kernel 5: flagged segmented sum with WARP-sized tiles ------------------------------------------
// Same flagged streams as kernel 3, but the unit of work is one warp = 256 consecutive entries, and
// warps never talk to each other: no shared memory, no __syncthreads.  The number of row-start flags
// before each warp's first entry comes from a table built at setup (warp_seg, 4 bytes per 256 entries),
// rows that straddle warps go through the carry arrays (one slot per warp) and spmv_fixup_kernel.
// Why: on a matrix without locality the x gathers queue in the SM's single in-order load pipeline at
// ~1 wavefront / clk, and every shared-memory round trip or barrier of kernel 3 waits behind that
// queue once more (ncu: short_scoreboard + barrier + mio stalls = 45 % in
// profiles/r01_k3_flag_ncu_summary.txt).  Here the only dependent round trips of a warp are
// streams -> gathers (-> seg_row -> y when the slab has empty rows / accumulates), and the 32 warps
// resident on an SM are 32 independent chains, so the load pipeline does not run dry.
constexpr int WK_ENTRIES = 256;  // per warp: 32 lanes x 8 entries
constexpr int WK_PER_TILE = (int)(B2_FLAG_TILE / WK_ENTRIES);

// ABL (measurement only, option "ablate"): 0 = the kernel; 1 = results kept in registers, no y / carry
// stores; 2 = also no segmented scan; 3 = no row logic at all (streams + gathers + multiply-add)
template <int DT, bool PERSIST, bool READ_Y, bool HAS_EMPTY, int ABL>
__global__ void __launch_bounds__(256, 4)
spmv_flag_warp_kernel(const uint32_t *__restrict__ colind, const uint64_t *__restrict__ vals,
                      const uint64_t *__restrict__ x, uint64_t *__restrict__ y,
                      const uint32_t *__restrict__ seg_row, uint32_t rows, const uint32_t *__restrict__ warp_seg,
                      uint32_t *__restrict__ carry_row, uint64_t *__restrict__ carry_val, uint32_t nwarps)
{
    constexpr int EPT = 8;
    const int lane = threadIdx.x & 31;
    const uint32_t gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (gw >= nwarps) return;
    const uint64_t pos = (uint64_t)gw * WK_ENTRIES + (uint64_t)lane * EPT;

    uint32_t c[EPT];
    uint64_t p[EPT];
    ld256_u32(colind + pos, c);
    ld256_u64(vals + pos, p);
    ld256_u64(vals + pos + 4, p + 4);
    const uint32_t wbase = __ldg(warp_seg + gw);
    const uint64_t pol_keep = policy_evict_last();
    uint64_t xv[EPT];
#pragma unroll
    for (int i = 0; i < EPT; i++) xv[i] = load_x<PERSIST>(x, c[i] & 0x7FFFFFFFu, pol_keep);

    if (ABL == 3) {
        uint64_t a = 0;
#pragma unroll
        for (int i = 0; i < EPT; i++) a = Semi<DT>::add(a, Semi<DT>::mul(p[i], xv[i]));
        if (a == 0x123456789abcdefull) y[0] = a + wbase;
        return;
    }
    uint32_t f = 0;
#pragma unroll
    for (int i = 0; i < EPT; i++) f |= (c[i] >> 31) << i;
    const uint32_t nfl = __popc(f);
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t lane_prefix = 0, warp_flags = 0;
#pragma unroll
    for (int bit = 0; (1 << bit) <= EPT; bit++) {
        const uint32_t bal_b = __ballot_sync(0xffffffffu, (nfl >> bit) & 1u);
        lane_prefix += __popc(bal_b & lt) << bit;
        warp_flags += __popc(bal_b) << bit;
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, nfl != 0);
    const uint32_t segs_before = wbase + lane_prefix;
    // the row my first flag closes and (accumulate / later slabs) its current y: issued while the
    // gathers are in flight
    uint32_t close_row = NO_ROW;
    uint64_t y_old = 0;
    if (nfl != 0 && segs_before != 0) {
        close_row = HAS_EMPTY ? __ldg(seg_row + segs_before - 1) : segs_before - 1;
        if (close_row >= rows) close_row = NO_ROW;
        if (READ_Y && close_row != NO_ROW) y_old = y[close_row];
    }

    uint64_t acc = 0, head = 0;
    uint32_t seen = 0;
#pragma unroll
    for (int i = 0; i < EPT; i++) {
        p[i] = Semi<DT>::mul(p[i], xv[i]);
        const uint32_t fl = (f >> i) & 1u;
        head = (fl & ~seen & 1u) ? acc : head;
        seen |= fl;
        acc = fl ? 0ull : acc;
        acc = Semi<DT>::add(acc, p[i]);
    }
    // segmented inclusive scan of the open tails: S = sum from the nearest flagged lane at or below me
    const uint32_t below = bal & ((2u << lane) - 1u);
    const int start = below ? 31 - __clz(below) : 0;
    uint64_t S = acc;
    uint64_t carry_in = 0;
    if (ABL < 2) {
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint64_t n = __shfl_up_sync(0xffffffffu, S, d);
            if (lane - d >= start) S = Semi<DT>::add(n, S);
        }
        carry_in = __shfl_up_sync(0xffffffffu, S, 1);
        if (lane == 0) carry_in = 0;
    }
    if (ABL >= 1) {
        uint64_t a = Semi<DT>::add(Semi<DT>::add(S, carry_in), Semi<DT>::add(head, y_old));
#pragma unroll
        for (int i = 0; i < EPT; i++) a ^= p[i];
        if (a == 0x123456789abcdefull) y[close_row & 1023] = a + start;
        return;
    }

    if (nfl != 0) {
        if (close_row != NO_ROW) {
            // when no lane below me has a flag this is only the warp's share of the row: the shares of
            // the preceding warps arrive through their carries
            uint64_t val = Semi<DT>::add(carry_in, head);
            if (READ_Y) val = Semi<DT>::add(y_old, val);
            y[close_row] = val;
        }
        if (nfl > 1) {
            uint64_t a = 0;
            uint32_t k = 0;
#pragma unroll
            for (int i = 0; i < EPT; i++) {
                if ((f >> i) & 1u) {
                    if (k != 0) store_segment<DT, READ_Y, HAS_EMPTY>(y, seg_row, rows, segs_before + k - 1, a);
                    k++;
                    a = 0;
                }
                a = Semi<DT>::add(a, p[i]);
            }
        }
    }
    if (lane == 31) {  // carry-out: the row left open at the end of this warp's entries
        const uint32_t seg_total = wbase + warp_flags;
        uint32_t row = NO_ROW;
        if (seg_total != 0) {
            row = HAS_EMPTY ? __ldg(seg_row + seg_total - 1) : seg_total - 1;
            if (row >= rows) row = NO_ROW;
        }
        carry_row[gw] = row;
        carry_val[gw] = S;
    }
}

// ---- hypersparse slabs: one thread per NON-EMPTY row --------------------------------------------
// A column slab of a matrix with far more columns than fit L2 (banded + random over 128M columns cut into
// 18 slabs) holds well under two entries per non-empty row.  The flagged kernel then has a row-start flag
// on almost every entry and falls into its per-flag path; here thread k simply owns the k-th non-empty row
// (seg_row, built at setup), walks its one or two entries in storage order and read-modify-writes y.
// All loads of neighbouring threads are neighbours (entries of consecutive non-empty rows are consecutive),
// y is touched in ascending order.  Summation order is the reference's, so this path is bit-exact in fp64.
// (A variant with four rows per thread and their loads issued together was slower: 2.30 vs 1.98 ms on the C5
// shard -- neighbouring lanes no longer touch neighbouring entries.  profiles/r01_c5_shard_sweep.txt)
template <int DT, bool PERSIST, bool READ_Y>
__global__ void __launch_bounds__(256)
spmv_dcsr_kernel(const uint32_t *__restrict__ rowptr, const uint32_t *__restrict__ colind, const uint64_t *__restrict__ vals,
                 const uint64_t *__restrict__ x, uint64_t *__restrict__ y, const uint32_t *__restrict__ seg_row,
                 uint32_t nseg, uint32_t colmask)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nseg) return;
    const uint32_t row = __ldg(seg_row + k);
    const uint32_t b = __ldg(rowptr + row), e = __ldg(rowptr + row + 1);
    uint64_t acc = READ_Y ? y[row] : 0ull;
    const uint64_t pol_keep = policy_evict_last();
    for (uint32_t p = b; p < e; p++) {
        const uint32_t c = __ldg(colind + p) & colmask;
        acc = Semi<DT>::add(acc, Semi<DT>::mul(__ldg(reinterpret_cast<const unsigned long long *>(vals + p)), load_x<PERSIST>(x, c, pol_keep)));
    }
    y[row] = acc;
}

// ---- setup of the flagged format (once per matrix) ---------------------------------------------
__global__ void nonempty_kernel(const uint32_t *__restrict__ rowptr, uint32_t rows, uint32_t *__restrict__ ne)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r <= rows) ne[r] = (r < rows && rowptr[r + 1] > rowptr[r]) ? 1u : 0u;
}
// segidx = exclusive scan of ne.  Flag each non-empty row's first entry, fill seg_row, flag the
// first pad entry (sentinel), terminate seg_row with NO_ROW.
__global__ void set_flags_kernel(const uint32_t *__restrict__ rowptr, const uint32_t *__restrict__ segidx, uint32_t rows,
                                 uint32_t *__restrict__ colind, uint32_t *__restrict__ seg_row)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < rows) {
        const uint32_t b = rowptr[r];
        if (rowptr[r + 1] > b) { colind[b] |= 0x80000000u; seg_row[segidx[r]] = r; }
    } else if (r == rows) {
        colind[rowptr[rows]] |= 0x80000000u;
        seg_row[segidx[rows]] = NO_ROW;
    }
}
// tile_seg[t] = flags strictly before entry slab_off + t*TILE
__global__ void tile_seg_kernel(const uint32_t *__restrict__ rowptr, const uint32_t *__restrict__ segidx, uint32_t rows,
                                uint32_t slab_off, uint32_t ntiles, uint32_t tile, uint32_t *__restrict__ tile_seg)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > ntiles) return;
    const uint64_t pos = (uint64_t)slab_off + (uint64_t)t * tile;
    uint32_t lo = 0, hi = rows + 1;  // first r in [0, rows] with rowptr[r] >= pos, else rows+1
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if ((uint64_t)rowptr[mid] < pos) lo = mid + 1; else hi = mid;
    }
    const uint32_t sentinel = pos > (uint64_t)rowptr[rows] ? 1u : 0u;
    tile_seg[t] = segidx[min(lo, rows)] + sentinel;
}

// Carries of rows that straddle tiles: the head of each run of equal carry rows adds the run
// (in tile order) to y.  Runs after spmv_merge_kernel on the same stream.
template <int DT>
__global__ void spmv_fixup_kernel(const uint32_t *__restrict__ carry_row, const uint64_t *__restrict__ carry_val,
                                  uint32_t ntiles, uint64_t *__restrict__ y)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    const uint32_t row = carry_row[t];
    if (row == NO_ROW) return;
    if (t > 0 && carry_row[t - 1] == row) return;  // not the head of its run
    uint64_t c = carry_val[t];
    for (uint32_t q = t + 1; q < ntiles && carry_row[q] == row; q++) c = Semi<DT>::add(c, carry_val[q]);
    y[row] = Semi<DT>::add(c, y[row]);
}

// The same for a launch cut into pieces of tiles (b200spmv_exec's download pipeline).  A row left open by tile t
// is stored by the tile that holds the next row start, so its carries may only be added once THAT tile has run:
// the run of carries that ends at tile e (carry_row[e + 1] differs) is added by the piece that contains tile e + 1.
// Thread e of [lo, hi) owns the runs that end at e; same order of additions as spmv_fixup_kernel.
template <int DT>
__global__ void spmv_fixup_range_kernel(const uint32_t *__restrict__ carry_row, const uint64_t *__restrict__ carry_val,
                                        uint32_t lo, uint32_t hi, uint32_t ntiles, uint64_t *__restrict__ y)
{
    const uint32_t e = lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= hi) return;
    const uint32_t row = carry_row[e];
    if (row == NO_ROW) return;
    if (e + 1 < ntiles && carry_row[e + 1] == row) return;  // not the end of its run
    uint32_t h = e;
    while (h > 0 && carry_row[h - 1] == row) h--;
    uint64_t c = carry_val[h];
    for (uint32_t q = h + 1; q <= e; q++) c = Semi<DT>::add(c, carry_val[q]);
    y[row] = Semi<DT>::add(c, y[row]);
}

// ---- strict order: one thread per row, reference summation order -------------------------
template <int DT, bool READ_Y>
__global__ void __launch_bounds__(256)
spmv_rowseq_kernel(const uint32_t *__restrict__ rowptr, const uint32_t *__restrict__ colind,
                   const uint64_t *__restrict__ vals, const uint64_t *__restrict__ x,
                   uint64_t *__restrict__ y, uint32_t rows, uint32_t colmask)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    uint64_t acc = READ_Y ? y[r] : 0ull;
    const uint32_t end = rowptr[r + 1];
    for (uint32_t p = rowptr[r]; p < end; p++)
        acc = Semi<DT>::add(acc, Semi<DT>::mul(vals[p], x[colind[p] & colmask]));  // mul then add, never FMA
    y[r] = acc;
}

}  // namespace

// ---- host side -----------------------------------------------------------------------------

// Merge-path tile tables for every slab.  Slab s owns tile-table entries
// [h_tile_base[s] + s, h_tile_base[s+1] + s] (ntiles_s + 1 of them) and carries [h_tile_base[s], ..).
int b2_build_tiles(b200spmv_ctx *ctx)
{
    DeviceCsr &m = ctx->csr;
    b2_free(m.tile_row); b2_free(m.tile_nz); b2_free(m.carry_row);
    { uint64_t *cv = (uint64_t *)m.carry_val; b2_free(cv); m.carry_val = nullptr; }
    std::vector<uint32_t> h_ptr(2 * (size_t)ctx->slabs);
    const size_t stride = ctx->rowptr_stride;
    uint32_t total = 0;
    for (int s = 0; s < ctx->slabs; s++) {
        uint32_t b = 0, e = 0;
        B2_CHECK(ctx, cudaMemcpyAsync(&b, m.rowptr + s * stride, 4, cudaMemcpyDeviceToHost, ctx->stream));
        B2_CHECK(ctx, cudaMemcpyAsync(&e, m.rowptr + s * stride + ctx->local_rows, 4, cudaMemcpyDeviceToHost, ctx->stream));
        B2_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        const uint64_t items = (uint64_t)ctx->local_rows + (e - b);
        ctx->h_tile_base[s] = total;
        total += b2_div_up(items, MP_TILE);
    }
    ctx->h_tile_base[ctx->slabs] = total;
    ctx->total_tiles = total;
    B2_TRY(b2_alloc(ctx, &m.tile_row, (size_t)total + ctx->slabs));
    B2_TRY(b2_alloc(ctx, &m.tile_nz, (size_t)total + ctx->slabs));
    B2_TRY(b2_alloc(ctx, &m.carry_row, (size_t)total + 1));
    uint64_t *cv = nullptr;
    B2_TRY(b2_alloc(ctx, &cv, (size_t)total + 1));
    m.carry_val = cv;
    for (int s = 0; s < ctx->slabs; s++) {
        const uint32_t nt = ctx->h_tile_base[s + 1] - ctx->h_tile_base[s];
        tile_search_kernel<<<b2_div_up((uint64_t)nt + 1, 256), 256, 0, ctx->stream>>>(
            m.rowptr + s * stride, ctx->local_rows, nt, m.tile_row + ctx->h_tile_base[s] + s,
            m.tile_nz + ctx->h_tile_base[s] + s);
        B2_CHECK(ctx, cudaGetLastError());
    }
    B2_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

// Flagged format for kernel 3 (needs bit 31 of the column ids free).
// longest row piece of a slab: the thread-per-row kernel is only chosen when no thread gets a long one
__global__ void max_rowlen_kernel(const uint32_t *__restrict__ rowptr, uint32_t rows, uint32_t *__restrict__ out)
{
    uint32_t v = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += (uint64_t)gridDim.x * blockDim.x)
        v = max(v, rowptr[i + 1] - rowptr[i]);
    v = __reduce_max_sync(0xFFFFFFFFu, v);
    if ((threadIdx.x & 31) == 0 && v > 0) atomicMax(out, v);
}

// Row below which everything is final once the tiles before tile T are done: the row that holds the entry just
// before T's first one may go on into T, every row before it cannot (tile_seg[T] = row starts before that entry).
__global__ void chunk_rows_kernel(const uint32_t *__restrict__ tile_seg, const uint32_t *__restrict__ seg_row, uint32_t ntiles,
                                  uint32_t rows, int chunks, uint32_t *__restrict__ out_tile, uint32_t *__restrict__ out_row)
{
    const int k = threadIdx.x;
    if (k > chunks) return;
    const uint32_t T = (uint32_t)((uint64_t)ntiles * k / chunks);
    uint32_t r = 0;
    if (k == chunks) r = rows;
    else if (T > 0) {
        const uint32_t c = tile_seg[T];
        r = c ? seg_row[c - 1] : 0;
        if (r > rows) r = rows;  // NO_ROW: only pads are left
    }
    out_tile[k] = T;
    out_row[k] = r;
}

int b2_build_flags(b200spmv_ctx *ctx)
{
    DeviceCsr &m = ctx->csr;
    cudaStream_t s = ctx->stream;
    b2_free(m.seg_row); b2_free(m.tile_seg); b2_free(m.warp_seg);
    ctx->flagged = false;
    ctx->colmask = 0xFFFFFFFFu;
    if (ctx->cols > 0x80000000u || ctx->local_rows == 0) return 0;  // kernels 1/2 only
    const uint32_t rows = ctx->local_rows;
    const size_t stride = ctx->rowptr_stride;
    uint32_t total = 0;
    for (int sl = 0; sl < ctx->slabs; sl++) {
        ctx->h_ftile_base[sl] = total;
        total += (uint32_t)((ctx->h_slab_off[sl + 1] - ctx->h_slab_off[sl]) / B2_FLAG_TILE);
    }
    ctx->h_ftile_base[ctx->slabs] = total;
    B2_TRY(b2_alloc(ctx, &m.seg_row, stride * ctx->slabs));
    B2_TRY(b2_alloc(ctx, &m.tile_seg, (size_t)total + ctx->slabs));
    B2_TRY(b2_alloc(ctx, &m.warp_seg, (size_t)total * WK_PER_TILE + ctx->slabs));
    if ((uint64_t)total * WK_PER_TILE > ctx->total_tiles) {  // the carry arrays serve every kernel: one slot per warp tile
        b2_free(m.carry_row);
        { uint64_t *cv = (uint64_t *)m.carry_val; b2_free(cv); m.carry_val = nullptr; }
        B2_TRY(b2_alloc(ctx, &m.carry_row, (size_t)total * WK_PER_TILE + 1));
        uint64_t *cv = nullptr;
        B2_TRY(b2_alloc(ctx, &cv, (size_t)total * WK_PER_TILE + 1));
        m.carry_val = cv;
    }
    uint32_t *segidx = nullptr;
    B2_TRY(b2_alloc(ctx, &segidx, (size_t)rows + 4));
    int rc = 0;
    for (int sl = 0; sl < ctx->slabs && rc == 0; sl++) {
        const uint32_t *rowptr = m.rowptr + sl * stride;
        nonempty_kernel<<<b2_div_up((uint64_t)rows + 1, 256), 256, 0, s>>>(rowptr, rows, segidx);
        if ((rc = b2_exclusive_scan_u32(ctx, segidx, (uint64_t)rows + 1, s))) break;
        set_flags_kernel<<<b2_div_up((uint64_t)rows + 1, 256), 256, 0, s>>>(rowptr, segidx, rows, m.colind, m.seg_row + sl * stride);
        const uint32_t nt = ctx->h_ftile_base[sl + 1] - ctx->h_ftile_base[sl];
        tile_seg_kernel<<<b2_div_up((uint64_t)nt + 1, 256), 256, 0, s>>>(rowptr, segidx, rows, (uint32_t)ctx->h_slab_off[sl], nt,
                                                                        B2_FLAG_TILE, m.tile_seg + ctx->h_ftile_base[sl] + sl);
        tile_seg_kernel<<<b2_div_up((uint64_t)nt * WK_PER_TILE + 1, 256), 256, 0, s>>>(
            rowptr, segidx, rows, (uint32_t)ctx->h_slab_off[sl], nt * WK_PER_TILE, WK_ENTRIES,
            m.warp_seg + (size_t)ctx->h_ftile_base[sl] * WK_PER_TILE + sl);
        cudaError_t e = cudaMemcpyAsync(&ctx->h_nseg[sl], segidx + rows, 4, cudaMemcpyDeviceToHost, s);
        // segidx[rows + 1] is free after the scan: scratch word for the longest row piece
        if (e == cudaSuccess) e = cudaMemsetAsync(segidx + rows + 1, 0, 4, s);
        max_rowlen_kernel<<<ctx->sms * 8, 256, 0, s>>>(rowptr, rows, segidx + rows + 1);
        if (e == cudaSuccess) e = cudaMemcpyAsync(&ctx->h_maxrow[sl], segidx + rows + 1, 4, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; }
    }
    ctx->chunks_ready = false;
    if (rc == 0) {  // download pipeline of b200spmv_exec: tile pieces of the last slab and the rows each one completes
        const int sl = ctx->slabs - 1;
        const uint32_t nt = ctx->h_ftile_base[sl + 1] - ctx->h_ftile_base[sl];
        uint32_t h[2 * (Y_CHUNKS + 1)];
        chunk_rows_kernel<<<1, 32, 0, s>>>(m.tile_seg + ctx->h_ftile_base[sl] + sl, m.seg_row + sl * stride, nt, rows, Y_CHUNKS,
                                           segidx, segidx + Y_CHUNKS + 1);
        cudaError_t e = cudaMemcpyAsync(h, segidx, sizeof h, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); ctx->cuda_error = (int)e; rc = B200SPMV_ECUDA; }
        else {
            bool ok = true;
            for (int k = 0; k <= Y_CHUNKS; k++) {
                ctx->h_chunk_tile[k] = h[k];
                ctx->h_chunk_row[k] = h[Y_CHUNKS + 1 + k];
                if (k && (ctx->h_chunk_row[k] < ctx->h_chunk_row[k - 1] || ctx->h_chunk_tile[k] < ctx->h_chunk_tile[k - 1])) ok = false;
            }
            ctx->chunks_ready = ok && ctx->h_chunk_row[0] == 0 && ctx->h_chunk_row[Y_CHUNKS] == rows && ctx->h_chunk_tile[Y_CHUNKS] == nt;
        }
    }
    cudaFree(segidx);
    if (rc) return rc;
    ctx->flagged = true;
    ctx->colmask = 0x7FFFFFFFu;
    return 0;
}

// Will b2_spmv_launch cut the last slab's kernel into row chunks (and record ev_rowchunk[])?  Kernel 3 on the
// flagged layout only: its tiles are consecutive pieces of the entry stream, hence of the rows.
static bool dcsr_slab(const b200spmv_ctx *ctx, int sl)
{
    const uint32_t nseg = ctx->h_nseg[sl];
    return ctx->opt_dcsr_x16 > 0 && nseg > 0 && (uint64_t)ctx->h_slab_nnz[sl] * 16 < (uint64_t)nseg * ctx->opt_dcsr_x16 &&
           ctx->h_maxrow[sl] <= (uint32_t)ctx->opt_dcsr_maxrow;
}
bool b2_last_slab_chunkable(const b200spmv_ctx *ctx, int flags)
{
    return ctx->chunks_ready && ctx->flagged && ctx->opt_kernel == 3 && !(flags & B200SPMV_STRICT_ORDER) &&
           ctx->order_n == 0 && ctx->local_rows > 0 && !dcsr_slab(ctx, ctx->slabs - 1);
}

static int set_persist_window(b200spmv_ctx *ctx, cudaStream_t s, const void *base, size_t bytes, bool on)
{
    cudaStreamAttrValue attr;
    memset(&attr, 0, sizeof attr);
    if (on && ctx->max_persist_bytes > 0 && bytes > 0) {
        int max_win = 0;
        cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, ctx->device);
        const size_t win = std::min(bytes, (size_t)max_win);
        size_t set_aside = ctx->max_persist_bytes;
        if (ctx->opt_persist_bytes > 0) set_aside = std::min(set_aside, (size_t)ctx->opt_persist_bytes);
        attr.accessPolicyWindow.base_ptr = const_cast<void *>(base);
        attr.accessPolicyWindow.num_bytes = win;
        attr.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)set_aside / (double)win);
        attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    } else {
        attr.accessPolicyWindow.num_bytes = 0;
        attr.accessPolicyWindow.hitRatio = 0.f;
        attr.accessPolicyWindow.hitProp = cudaAccessPropertyNormal;
        attr.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
    }
    B2_CHECK(ctx, cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &attr));
    return 0;
}

// The slab's slice of x, streamed into L2 by the TMA unit ahead of the gathers (cp.async.bulk.prefetch.L2 ->
// UBLKPF, nothing returns to the SM): the first gather of every sector is then an L2 hit instead of a random
// 32-byte DRAM read (42-65 G/s, profiles/r01_gather_paths.txt).  One 16 KB request per thread.
constexpr uint32_t XPF_CHUNK = 16384;
__global__ void x_prefetch_kernel(const char *base, uint64_t bytes)
{
    const uint64_t off = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * XPF_CHUNK;
    if (off < bytes) {
        const uint32_t n = (uint32_t)min((uint64_t)XPF_CHUNK, bytes - off);
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(base + off), "r"(n) : "memory");
    }
}

template <int DT>
static int launch_dtype(b200spmv_ctx *ctx, int flags, cudaStream_t s)
{
    const DeviceCsr &m = ctx->csr;
    const bool strict = flags & B200SPMV_STRICT_ORDER;
    const bool persist = !(flags & B200SPMV_NO_PERSIST);
    const bool accumulate = flags & B200SPMV_ACCUMULATE;
    const uint64_t *x = (const uint64_t *)ctx->d_x;
    uint64_t *y = (uint64_t *)ctx->d_y + ctx->row_begin;
    const uint64_t *vals = (const uint64_t *)m.vals;
    const size_t stride = ctx->rowptr_stride;
    const uint32_t rows = ctx->local_rows;
    if (rows == 0) return 0;
    // slab order: column order, unless the iterated SpMV has put this rank's own block of x first (common.cuh)
    const bool ordered = ctx->order_n == ctx->slabs;
    for (int pos = 0; pos < ctx->slabs; pos++) {
        const int sl = ordered ? ctx->h_order[pos] : pos;
        if (ordered && ctx->slab_wait[pos]) B2_CHECK(ctx, cudaStreamWaitEvent(s, ctx->slab_wait[pos], 0));
        const uint32_t *rowptr = m.rowptr + sl * stride;
        const bool read_y = accumulate || pos > 0;
        if (ctx->opt_prefetch_x && ctx->slabs > 1) {
            // whole 16-byte units around the slice (every device array is padded by 256 bytes)
            const uint64_t b0 = (uint64_t)ctx->h_slab_col[sl] * 8 & ~15ull, b1 = ((uint64_t)ctx->h_slab_col[sl + 1] * 8 + 15) & ~15ull;
            if (b1 > b0 && (double)(b1 - b0) <= 0.6 * (double)ctx->l2_bytes) {
                const uint64_t nthr = (b1 - b0 + XPF_CHUNK - 1) / XPF_CHUNK;
                x_prefetch_kernel<<<(unsigned)((nthr + 127) / 128), 128, 0, s>>>((const char *)x + b0, b1 - b0);
                ctx->launches += 1;
            }
        }
        if (ctx->wait_slab_events) B2_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_slab[sl], 0));
        if (!strict && ctx->opt_kernel >= 3 && ctx->flagged) {
            const uint32_t tb = ctx->h_ftile_base[sl];
            const uint32_t nt = ctx->h_ftile_base[sl + 1] - tb;
            const bool has_empty = ctx->h_nseg[sl] < rows;
            if (!read_y && has_empty) B2_CHECK(ctx, cudaMemsetAsync(y, 0, (size_t)rows * 8, s));  // rows no flag will touch
            const uint32_t *ci = m.colind + ctx->h_slab_off[sl];
            const uint64_t *va = vals + ctx->h_slab_off[sl];
            const uint32_t *segrow = m.seg_row + sl * stride, *tseg = m.tile_seg + tb + sl;
            uint32_t *crow = m.carry_row + tb;
            uint64_t *cval = (uint64_t *)m.carry_val + tb;
            const bool prof = ctx->opt_profile && ctx->prof_n + 2 <= 128;
            if (prof) {
                if (!ctx->prof_ev[ctx->prof_n]) { cudaEventCreate(&ctx->prof_ev[ctx->prof_n]); cudaEventCreate(&ctx->prof_ev[ctx->prof_n + 1]); }
                cudaEventRecord(ctx->prof_ev[ctx->prof_n], s);
            }
            // hypersparse slab (fewer than opt_dcsr_x16/16 entries per non-empty row and no long row: on a skewed
            // matrix one thread with a few thousand entries holds the whole launch up): thread per non-empty row
            const uint32_t nseg = ctx->h_nseg[sl];
            if (dcsr_slab(ctx, sl)) {
                const unsigned grid = b2_div_up(nseg, 256);
                if (persist) {
                    if (read_y) spmv_dcsr_kernel<DT, true, true><<<grid, 256, 0, s>>>(rowptr, m.colind, vals, x, y, segrow, nseg, ctx->colmask);
                    else spmv_dcsr_kernel<DT, true, false><<<grid, 256, 0, s>>>(rowptr, m.colind, vals, x, y, segrow, nseg, ctx->colmask);
                } else {
                    if (read_y) spmv_dcsr_kernel<DT, false, true><<<grid, 256, 0, s>>>(rowptr, m.colind, vals, x, y, segrow, nseg, ctx->colmask);
                    else spmv_dcsr_kernel<DT, false, false><<<grid, 256, 0, s>>>(rowptr, m.colind, vals, x, y, segrow, nseg, ctx->colmask);
                }
                if (prof) { cudaEventRecord(ctx->prof_ev[ctx->prof_n + 1], s); ctx->prof_n += 2; }
                ctx->launches += 1;
                B2_CHECK(ctx, cudaGetLastError());
                continue;
            }
#define B2_LAUNCH_FLAG(P, R, E)                                                                                     \
    do {                                                                                                           \
        if (ctx->opt_ept == 16)                                                                                    \
            spmv_flag_kernel<DT, P, R, E, 16><<<pn, B2_FLAG_TILE / 16, 0, s>>>(pci, pva, x, y, segrow, rows, ptseg, pcrow, pcval, (uint32_t)ctx->opt_prefetch_tiles); \
        else                                                                                                       \
            spmv_flag_kernel<DT, P, R, E, 8><<<pn, B2_FLAG_TILE / 8, 0, s>>>(pci, pva, x, y, segrow, rows, ptseg, pcrow, pcval, (uint32_t)ctx->opt_prefetch_tiles);  \
    } while (0)
#define B2_LAUNCH_RING(P, R, E)                                                                                     \
    do {                                                                                                           \
        auto kfn = spmv_flag_ring_kernel<DT, P, R, E>;                                                             \
        static bool attr_set = false;                                                                              \
        if (!attr_set) {                                                                                           \
            B2_CHECK(ctx, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PK_SMEM_BYTES)); \
            B2_CHECK(ctx, cudaFuncSetAttribute(kfn, cudaFuncAttributePreferredSharedMemoryCarveout, 100));         \
            attr_set = true;                                                                                       \
        }                                                                                                          \
        const uint32_t grid = std::min<uint32_t>(nt, (uint32_t)ctx->sms * (uint32_t)ctx->opt_ctas_per_sm);         \
        kfn<<<grid, B2_FLAG_TILE / 8, PK_SMEM_BYTES, s>>>(ci, va, x, y, segrow, rows, tseg, crow, cval, nt);       \
    } while (0)
#define B2_LAUNCH_WARP(P, R, E)                                                                                     \
    do {                                                                                                           \
        if (ctx->opt_ablate == 1)                                                                                  \
            spmv_flag_warp_kernel<DT, P, R, E, 1><<<b2_div_up(nw, 8), 256, 0, s>>>(ci, va, x, y, segrow, rows, wseg, wcrow, wcval, nw); \
        else if (ctx->opt_ablate == 2)                                                                             \
            spmv_flag_warp_kernel<DT, P, R, E, 2><<<b2_div_up(nw, 8), 256, 0, s>>>(ci, va, x, y, segrow, rows, wseg, wcrow, wcval, nw); \
        else if (ctx->opt_ablate == 3)                                                                             \
            spmv_flag_warp_kernel<DT, P, R, E, 3><<<b2_div_up(nw, 8), 256, 0, s>>>(ci, va, x, y, segrow, rows, wseg, wcrow, wcval, nw); \
        else                                                                                                       \
            spmv_flag_warp_kernel<DT, P, R, E, 0><<<b2_div_up(nw, 8), 256, 0, s>>>(ci, va, x, y, segrow, rows, wseg, wcrow, wcval, nw); \
    } while (0)
            const uint32_t nw = nt * WK_PER_TILE;
            const uint32_t *wseg = m.warp_seg + (size_t)tb * WK_PER_TILE + sl;
            uint32_t *wcrow = m.carry_row + (size_t)tb * WK_PER_TILE;
            uint64_t *wcval = (uint64_t *)m.carry_val + (size_t)tb * WK_PER_TILE;
            if (ctx->opt_kernel == 5) {
                if (persist) {
                    if (read_y) { if (has_empty) B2_LAUNCH_WARP(true, true, true); else B2_LAUNCH_WARP(true, true, false); }
                    else { if (has_empty) B2_LAUNCH_WARP(true, false, true); else B2_LAUNCH_WARP(true, false, false); }
                } else {
                    if (read_y) { if (has_empty) B2_LAUNCH_WARP(false, true, true); else B2_LAUNCH_WARP(false, true, false); }
                    else { if (has_empty) B2_LAUNCH_WARP(false, false, true); else B2_LAUNCH_WARP(false, false, false); }
                }
            } else if (ctx->opt_kernel == 4) {
                if (persist) {
                    if (read_y) { if (has_empty) B2_LAUNCH_RING(true, true, true); else B2_LAUNCH_RING(true, true, false); }
                    else { if (has_empty) B2_LAUNCH_RING(true, false, true); else B2_LAUNCH_RING(true, false, false); }
                } else {
                    if (read_y) { if (has_empty) B2_LAUNCH_RING(false, true, true); else B2_LAUNCH_RING(false, true, false); }
                    else { if (has_empty) B2_LAUNCH_RING(false, false, true); else B2_LAUNCH_RING(false, false, false); }
                }
            } else {
                // kernel 3.  For b200spmv_exec the last slab goes out in Y_CHUNKS pieces of consecutive tiles, each
                // followed by its own fix-up and an event: the rows a piece completes can leave for the host while
                // the pieces after it still run (common.cuh, h_chunk_row).
                const bool chunked = ctx->want_row_chunks && pos == ctx->slabs - 1 && b2_last_slab_chunkable(ctx, flags);
                const int pieces = chunked ? Y_CHUNKS : 1;
                for (int k = 0; k < pieces; k++) {
                    const uint32_t t0 = chunked ? ctx->h_chunk_tile[k] : 0, t1 = chunked ? ctx->h_chunk_tile[k + 1] : nt;
                    const uint32_t pn = t1 - t0;
                    if (pn > 0) {
                        const uint32_t *pci = ci + (size_t)t0 * B2_FLAG_TILE, *ptseg = tseg + t0;
                        const uint64_t *pva = va + (size_t)t0 * B2_FLAG_TILE;
                        uint32_t *pcrow = crow + t0;
                        uint64_t *pcval = cval + t0;
                        if (persist) {
                            if (read_y) { if (has_empty) B2_LAUNCH_FLAG(true, true, true); else B2_LAUNCH_FLAG(true, true, false); }
                            else { if (has_empty) B2_LAUNCH_FLAG(true, false, true); else B2_LAUNCH_FLAG(true, false, false); }
                        } else {
                            if (read_y) { if (has_empty) B2_LAUNCH_FLAG(false, true, true); else B2_LAUNCH_FLAG(false, true, false); }
                            else { if (has_empty) B2_LAUNCH_FLAG(false, false, true); else B2_LAUNCH_FLAG(false, false, false); }
                        }
                        ctx->launches += 1;
                    }
                    if (k == pieces - 1 && prof) { cudaEventRecord(ctx->prof_ev[ctx->prof_n + 1], s); ctx->prof_n += 2; }
                    if (!chunked) {
                        spmv_fixup_kernel<DT><<<b2_div_up(nt, 256), 256, 0, s>>>(crow, cval, nt, y);
                        ctx->launches += 1;
                    } else {
                        // carries whose run ends before the last tile this piece has run (the last piece: all that is left)
                        const uint32_t lo = k == 0 ? 0 : std::max(t0, 1u) - 1, hi = k == pieces - 1 ? nt : std::max(t1, 1u) - 1;
                        if (hi > lo) {
                            spmv_fixup_range_kernel<DT><<<b2_div_up(hi - lo, 256), 256, 0, s>>>(crow, cval, lo, hi, nt, y);
                            ctx->launches += 1;
                        }
                        B2_CHECK(ctx, cudaEventRecord(ctx->ev_rowchunk[k], s));
                    }
                }
                ctx->did_row_chunks = ctx->did_row_chunks || chunked;
            }
#undef B2_LAUNCH_FLAG
#undef B2_LAUNCH_RING
#undef B2_LAUNCH_WARP
            if (ctx->opt_kernel == 5 || ctx->opt_kernel == 4) {
                if (prof) { cudaEventRecord(ctx->prof_ev[ctx->prof_n + 1], s); ctx->prof_n += 2; }
                if (ctx->opt_kernel == 5) spmv_fixup_kernel<DT><<<b2_div_up(nw, 256), 256, 0, s>>>(wcrow, wcval, nw, y);
                else spmv_fixup_kernel<DT><<<b2_div_up(nt, 256), 256, 0, s>>>(crow, cval, nt, y);
                ctx->launches += 2;
            }
        } else if (strict) {
            const unsigned grid = b2_div_up(rows, 256);
            if (read_y) spmv_rowseq_kernel<DT, true><<<grid, 256, 0, s>>>(rowptr, m.colind, vals, x, y, rows, ctx->colmask);
            else spmv_rowseq_kernel<DT, false><<<grid, 256, 0, s>>>(rowptr, m.colind, vals, x, y, rows, ctx->colmask);
            ctx->launches += 1;
        } else {
            const uint32_t tb = ctx->h_tile_base[sl];
            const uint32_t nt = ctx->h_tile_base[sl + 1] - tb;
            if (nt == 0) continue;
            const uint32_t *trow = m.tile_row + tb + sl, *tnz = m.tile_nz + tb + sl;
            uint32_t *crow = m.carry_row + tb;
            uint64_t *cval = (uint64_t *)m.carry_val + tb;
#define B2_LAUNCH_MERGE(P, R)                                                                                      \
    do {                                                                                                           \
        if (ctx->opt_kernel == 1)                                                                                  \
            spmv_merge_kernel<DT, P, R><<<nt, MP_THREADS, 0, s>>>(rowptr, m.colind, vals, x, y, rows, trow, tnz, crow, cval, ctx->colmask); \
        else                                                                                                       \
            spmv_merge_tma_kernel<DT, P, R><<<nt, MP_THREADS, 0, s>>>(rowptr, m.colind, vals, x, y, rows, trow, tnz, crow, cval, ctx->colmask); \
    } while (0)
            const bool prof = ctx->opt_profile && ctx->prof_n + 2 <= 128;
            if (prof) {
                if (!ctx->prof_ev[ctx->prof_n]) { cudaEventCreate(&ctx->prof_ev[ctx->prof_n]); cudaEventCreate(&ctx->prof_ev[ctx->prof_n + 1]); }
                cudaEventRecord(ctx->prof_ev[ctx->prof_n], s);
            }
            if (persist) { if (read_y) B2_LAUNCH_MERGE(true, true); else B2_LAUNCH_MERGE(true, false); }
            else { if (read_y) B2_LAUNCH_MERGE(false, true); else B2_LAUNCH_MERGE(false, false); }
#undef B2_LAUNCH_MERGE
            if (prof) { cudaEventRecord(ctx->prof_ev[ctx->prof_n + 1], s); ctx->prof_n += 2; }
            spmv_fixup_kernel<DT><<<b2_div_up(nt, 256), 256, 0, s>>>(crow, cval, nt, y);
            ctx->launches += 2;
        }
        B2_CHECK(ctx, cudaGetLastError());
    }
    ctx->last_mode = strict ? 0 : ((ctx->opt_kernel >= 3 && !ctx->flagged) ? 2 : ctx->opt_kernel);
    return 0;
}

// Fold the recorded merge-kernel event pairs into merge_kernel_ns (call after a sync).
void b2_profile_collect(b200spmv_ctx *ctx)
{
    for (int i = 0; i + 1 < ctx->prof_n; i += 2) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ctx->prof_ev[i], ctx->prof_ev[i + 1]) == cudaSuccess) {
            ctx->merge_kernel_ns += (uint64_t)((double)ms * 1e6);
            ctx->merge_kernel_launches++;
            ctx->prof_slab_ns[(i / 2) % ctx->slabs] += (uint64_t)((double)ms * 1e6);  // collected at step boundaries
        }
    }
    ctx->prof_n = 0;
}

// Everything a launch needs that is not a stream operation (safe to call before a graph capture).
// The access-policy window is a stream attribute: set when the mode changes, not per launch.  It covers the
// whole of x up to the device's window limit; which slice is hot is decided by the slab being processed,
// the window only marks x as the data to keep (profiles/r01_gather_paths.txt, A/B).
int b2_spmv_prepare(b200spmv_ctx *ctx, int flags, cudaStream_t s)
{
    const bool persist = !(flags & B200SPMV_NO_PERSIST);
    if (persist != ctx->persist_set) {
        B2_TRY(set_persist_window(ctx, s, ctx->d_x, (size_t)ctx->cols * 8, persist));
        ctx->persist_set = persist;
    }
    static bool carveout_set = false;
    if (!carveout_set) {
#define B2_CARVE(DT) \
    cudaFuncSetAttribute(spmv_merge_kernel<DT, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100); \
    cudaFuncSetAttribute(spmv_merge_kernel<DT, true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100); \
    cudaFuncSetAttribute(spmv_merge_kernel<DT, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100); \
    cudaFuncSetAttribute(spmv_merge_kernel<DT, false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100); \
    cudaFuncSetAttribute(spmv_merge_tma_kernel<DT, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100); \
    cudaFuncSetAttribute(spmv_merge_tma_kernel<DT, true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100); \
    cudaFuncSetAttribute(spmv_merge_tma_kernel<DT, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100); \
    cudaFuncSetAttribute(spmv_merge_tma_kernel<DT, false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        B2_CARVE(B200SPMV_F64)
        B2_CARVE(B200SPMV_U64)
#undef B2_CARVE
        carveout_set = true;
    }
    return 0;
}

int b2_spmv_launch(b200spmv_ctx *ctx, int flags, cudaStream_t s)
{
    B2_TRY(b2_spmv_prepare(ctx, flags, s));
    return ctx->dtype == B200SPMV_F64 ? launch_dtype<B200SPMV_F64>(ctx, flags, s)
                                      : launch_dtype<B200SPMV_U64>(ctx, flags, s);
}
