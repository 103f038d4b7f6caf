// spmv.cu -- the hot path: y (+)= A*x on the device-resident CSR (software/SoftwareSpMV.cpp:59-64
// after the CSC->CSR transpose).
//
//   spmv_flag_kernel (3, default) : flagged segmented sum, one 2048-entry tile per CTA.  Row-start flags in bit 31
//       of the column ids replace the merge-path search; exact nnz split (8 consecutive entries per thread, 256-bit
//       loads), warp-shuffle segmented scan, tiles stitched by spmv_fixup_kernel.  Its tiles are consecutive
//       pieces of the rows, so b200spmv_exec launches the last slab in pieces and sends finished rows to the host
//       while later pieces run.
//   spmv_stream_kernel (6) : the same arithmetic by PERSISTENT, autonomous warps.  Every warp owns a contiguous
//       chunk of the slab's entry stream; the open row travels in a register from step to step, the rows a step
//       closes are staged in shared memory and written out coalesced, row ids and old y are requested one step
//       ahead, one carry per chunk.  Round-2 experiment against the "second limiter" of VERDICT round 1: it ties
//       kernel 3 on gather-bound matrices (195 vs 196 GNNZ/s on C3), wins by 3-6 % where HBM binds (pure band:
//       420 vs 407) and loses on skewed rows (R-MAT: static chunks).  Selectable (option kernel = 6); DESIGN.md 3.
//   spmv_dcsr_kernel : hypersparse slabs, one thread per non-empty row (storage order: bit-exact in fp64).
//   spmv_rowseq_kernel : one thread per row, entries in CSR (= reference) order, separate
//       multiply and add: fp64 results memcmp-equal to SoftwareSpMV (strict mode).
//
// Column slabs: with S > 1 the CSR is stored slab-major and the kernels run slab by slab, so
// the x slice gathered at any time fits L2 (the "vector cache").
// The merge-path kernels (1, 2) and the TMA-ring kernel (4) of round 1 lost to kernel 3 by 2.5-2.9x and were
// removed; kernel 5 (warp tiles, one carry per 256 entries) is subsumed by kernel 6.  DESIGN.md 3.4 keeps
// their numbers.
#include <algorithm>
#include <vector>

#include "common.cuh"
#include <chrono>

namespace {

constexpr uint32_t NO_ROW = 0xFFFFFFFFu;

template <bool PERSIST>
__device__ __forceinline__ uint64_t load_x(const uint64_t *__restrict__ x, uint32_t col, uint64_t pol)
{
    if (PERSIST) return ld_gather_u64(x + col, pol);
    return __ldg(reinterpret_cast<const unsigned long long *>(x + col));
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

// y[row] (+)= v without reading y into the SM: one reduction request (RED.ADD.F64 / .64) instead of a load and a
// store.  Kernel 3 closes every row at most once per launch (the carries are added by the fix-up launch), so the
// result is the same y_old + v, bit for bit, as the load-add-store it replaces -- IEEE round-to-nearest on the
// L2's adder as on the SM's -- and the summation order stays deterministic.
#ifndef B2_Y_RED
#define B2_Y_RED 1  // 0: load-add-store (measurement builds: make variant NAME=nored DEFS=-DB2_Y_RED=0)
#endif
template <int DT>
__device__ __forceinline__ void y_reduce(uint64_t *addr, uint64_t v)
{
    if (DT == B200SPMV_F64) atomicAdd(reinterpret_cast<double *>(addr), __longlong_as_double((long long)v));
    else atomicAdd(reinterpret_cast<unsigned long long *>(addr), (unsigned long long)v);
}
template <int DT, bool READ_Y, bool HAS_EMPTY>
__device__ __forceinline__ void store_segment(uint64_t *__restrict__ y, const uint32_t *__restrict__ seg_row,
                                              uint32_t rows, uint32_t seg, uint64_t v)
{
    const uint32_t row = HAS_EMPTY ? __ldg(seg_row + seg) : seg;
    if (row < rows) {
        if (READ_Y && B2_Y_RED) { y_reduce<DT>(y + row, v); return; }
        if (READ_Y) v = Semi<DT>::add(y[row], v);
        y[row] = v;
    }
}
#ifndef B2_K3_CTAS
#define B2_K3_CTAS 4  // CTAs per SM the register budget is cut for (measurement builds: make variant DEFS=-DB2_K3_CTAS=5)
#endif
// EPT = consecutive entries per thread (8 or 16); the tile is always B2_FLAG_TILE entries, so the
// CTA has B2_FLAG_TILE / EPT threads.
template <int DT, bool PERSIST, bool READ_Y, bool HAS_EMPTY, int EPT>
__global__ void __launch_bounds__(B2_FLAG_TILE / EPT, B2_K3_CTAS)
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
        if (READ_Y && !B2_Y_RED && close_row != NO_ROW) y_old = y[close_row];
    }
    // a second flag in my entries closes a row that lies wholly inside them: its row id and y are fetched
    // now as well (rows shorter than 8 entries are common once a matrix is cut into column slabs)
    uint32_t row2 = NO_ROW;
    uint64_t y_old2 = 0;
    if (nfl > 1) {
        row2 = HAS_EMPTY ? __ldg(seg_row + segs_before) : segs_before;
        if (row2 >= rows) row2 = NO_ROW;
        if (READ_Y && !B2_Y_RED && row2 != NO_ROW) y_old2 = y[row2];
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
            if (READ_Y && B2_Y_RED) y_reduce<DT>(y + close_row, val);
            else y[close_row] = READ_Y ? Semi<DT>::add(y_old, val) : val;
        }
        if (nfl > 1 && row2 != NO_ROW) {
            if (READ_Y && B2_Y_RED) y_reduce<DT>(y + row2, mid);
            else y[row2] = READ_Y ? Semi<DT>::add(y_old2, mid) : mid;
        }
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
// ---- kernel 6: flagged segmented sum by persistent, autonomous warps ---------------------------
// Same flagged format and the same per-thread arithmetic as kernel 3 (8 consecutive entries per thread,
// 256 per warp step), but:
//   * a warp owns a CONTIGUOUS chunk of the slab's entry stream and loops over it, so the row left open at
//     the end of a step stays in a register (wcarry) and the flag count before the next step is a running
//     sum: no tile_seg / warp_seg lookups after the first, and one carry per CHUNK for the fix-up pass
//     instead of one per tile;
//   * the 12 B/entry streams of step i+1 are requested before the x gathers of step i are issued (register
//     double buffer; optionally the TMA unit also pulls step i+pf into L2), so every warp always has 3 KB of
//     HBM reads in flight.  A non-persistent CTA spends most of its life NOT loading streams (launch, gather
//     round trip, row logic, y round trip, exit), which is what held kernel 3 near 200 GNNZ/s whatever the
//     gather locality;
//   * no block barrier: the warps of an SM are independent chains;
//   * the sums of the rows a step closes go through a warp-private strip of shared memory to the lanes that
//     write them out, so y is read and written with coalesced requests (see stream_step).
constexpr int SK_THREADS = 256;
constexpr int WK_ENTRIES = 256;  // per warp step: 32 lanes x 8 entries
constexpr int WK_PER_TILE = (int)(B2_FLAG_TILE / WK_ENTRIES);

#ifndef B200SPMV_ABLATE
#define B200SPMV_ABLATE 0  // measurement builds only (make ablate): 1 no y stores, 3 no row logic, 4 no gathers either
#endif

__device__ __forceinline__ void stream_load(uint32_t (&c)[8], uint64_t (&p)[8], const uint32_t *cp, const uint64_t *vp)
{
    ld256_u32(cp, c);
    ld256_u64(vp, p);
    ld256_u64(vp + 4, p + 4);
}

// Row id of the segment (non-empty row) with index seg; seg may run past the slab's last segment (prefetch for a
// step that closes fewer rows): the table has rows + 1 entries per slab, anything past it reads as "no row".
template <bool HAS_EMPTY>
__device__ __forceinline__ uint32_t segment_row(const uint32_t *__restrict__ seg_row, uint32_t rows, uint32_t seg)
{
    if (!HAS_EMPTY) return seg;
    return seg <= rows ? __ldg(seg_row + seg) : NO_ROW;
}

// One warp step: 256 consecutive entries, c / p = this lane's 8 column ids (flag in bit 31) and values.
// segs_base = row-start flags before the step's first entry, wcarry = partial sum of the row open on entry;
// both are advanced to the end of the step.
//
// Every flag closes the row before it.  The sums of the rows a step closes are collected in a warp-private
// shared-memory strip sv[] -- slot j = j-th row closed by the step, i.e. segment segs_base - 1 + j -- and written
// out by lane j: consecutive segments are (nearly) consecutive rows, so y is read and written with ONE coalesced
// load and ONE coalesced store per 32 closed rows instead of one request per row.
// The step is one stage of a software pipeline over the warp's chunk: row_n / y_n arrive holding the row id and
// the old y of slot `lane` of THIS step (requested a whole step ago), and leave holding those of the next step
// -- its first segment is known as soon as this step's flags are counted.  ncu on the version that looked the row
// ids up inside the step: 20 % of all warp time waiting for the row-id table and as much again for y, with the
// SM's L2 request port -- the real limit of a gather-bound matrix -- only 71 % busy.
template <int DT, bool PERSIST, bool READ_Y, bool HAS_EMPTY>
__device__ __forceinline__ void stream_step(const uint32_t (&c)[8], uint64_t (&p)[8], const uint64_t *__restrict__ x,
                                            uint64_t *__restrict__ y, const uint32_t *__restrict__ seg_row, uint32_t rows,
                                            int lane, uint32_t lt, uint64_t pol_keep, uint64_t *__restrict__ sv,
                                            uint32_t &segs_base, uint64_t &wcarry, uint32_t &row_n, uint64_t &y_n)
{
    constexpr int ABL = B200SPMV_ABLATE;
    uint64_t xv[8];
    if (ABL >= 3) {
#pragma unroll
        for (int i = 0; i < 8; i++) xv[i] = ABL >= 4 ? (uint64_t)c[i] : load_x<PERSIST>(x, c[i] & 0x7FFFFFFFu, pol_keep);
        uint64_t a = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) a = Semi<DT>::add(a, Semi<DT>::mul(p[i], xv[i]));
        wcarry ^= a;
        return;
    }
    uint32_t f = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) f |= (c[i] >> 31) << i;
    const uint32_t nfl = __popc(f);
    // flags in lower lanes: sum over the bits of nfl of ballots (no dependent shuffle chain)
    uint32_t lane_prefix = 0, warp_flags = 0;
#pragma unroll
    for (int bit = 0; bit < 4; bit++) {
        const uint32_t bal_b = __ballot_sync(0xffffffffu, (nfl >> bit) & 1u);
        lane_prefix += __popc(bal_b & lt) << bit;
        warp_flags += __popc(bal_b) << bit;
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, nfl != 0);
    // next step's slot `lane`: segment (segs_base + warp_flags) - 1 + lane.  Its row id is requested now and its y
    // at the end of this step; both are consumed at the end of the next one.
    const uint32_t next_seg = segs_base + warp_flags + (uint32_t)lane - 1u;
    const uint32_t row_next = next_seg != 0xFFFFFFFFu ? segment_row<HAS_EMPTY>(seg_row, rows, next_seg) : NO_ROW;
#pragma unroll
    for (int i = 0; i < 8; i++) xv[i] = load_x<PERSIST>(x, c[i] & 0x7FFFFFFFu, pol_keep);
    // write-out duty of this lane: the lane-th row the step closes
    const uint32_t my_row = ((uint32_t)lane < warp_flags && row_n < rows) ? row_n : NO_ROW;
    const uint64_t y_old = y_n;

    // branch-free sequential pass: head = sum before my first flag, acc = sum after my last flag; the rows that
    // begin and end inside my 8 entries go straight to their slots
    const uint32_t f1 = f & (0u - f);
    uint64_t *slot = sv + lane_prefix;  // slot of the row my next flag closes
    uint64_t acc = 0, head = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        p[i] = Semi<DT>::mul(p[i], xv[i]);
        const bool fl = (f >> i) & 1u, first = (f1 >> i) & 1u;
        head = first ? acc : head;
        if (fl && !first) *slot = acc;
        slot += fl;
        acc = fl ? 0ull : acc;
        acc = Semi<DT>::add(acc, p[i]);
    }
    // segmented inclusive scan of the open tails over the warp: S = sum of tails from the nearest flagged lane at
    // or below me (or lane 0) to me.  Steps longer than the warp's longest open run change nothing: skipped.
    const uint32_t below = bal & ((2u << lane) - 1u);
    const int start = below ? 31 - __clz(below) : 0;
    uint64_t S = acc;
    const int span = __reduce_max_sync(0xffffffffu, lane - start);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        if (d > span) break;
        const uint64_t n = __shfl_up_sync(0xffffffffu, S, d);
        if (lane - d >= start) S = Semi<DT>::add(n, S);
    }
    uint64_t prev = __shfl_up_sync(0xffffffffu, S, 1);  // open partial that ends at lane - 1
    if (lane == 0) prev = 0;
    const uint64_t s31 = __shfl_sync(0xffffffffu, S, 31);
    if (nfl != 0) {
        // my first flag closes the row open before my entries; with no flagged lane below me that row was open
        // when the step began and its earlier part is wcarry
        const uint64_t val = (bal & lt) == 0 ? Semi<DT>::add(wcarry, prev) : prev;
        sv[lane_prefix] = Semi<DT>::add(val, head);
    }
    __syncwarp();
    if (my_row != NO_ROW) {
        const uint64_t v = READ_Y ? Semi<DT>::add(y_old, sv[lane]) : sv[lane];
        if (ABL >= 1) wcarry ^= v; else y[my_row] = v;
    }
    for (uint32_t j = 32u + (uint32_t)lane; j < warp_flags; j += 32u) {  // short rows: more than 32 closed in one step
        const uint32_t row = segment_row<HAS_EMPTY>(seg_row, rows, segs_base + j - 1u);
        if (row < rows) {
            const uint64_t v = READ_Y ? Semi<DT>::add(y[row], sv[j]) : sv[j];
            if (ABL >= 1) wcarry ^= v; else y[row] = v;
        }
    }
    __syncwarp();  // sv is rewritten by the next step
    wcarry = bal ? s31 : Semi<DT>::add(wcarry, s31);
    segs_base += warp_flags;
    row_n = row_next;
    if (READ_Y) y_n = row_next < rows ? y[row_next] : 0ull;
}

// nwt = warp steps (256 entries) of the slab, cut into nchunks chunks of `chunk` steps; warp w of the grid takes
// chunks w, w + warps, ...  DB: register double buffer for the streams.  pf_dist: L2 bulk prefetch distance in
// steps (0 = off).
template <int DT, bool PERSIST, bool READ_Y, bool HAS_EMPTY, bool DB>
__global__ void __launch_bounds__(SK_THREADS, DB ? 2 : 3)
spmv_stream_kernel(const uint32_t *__restrict__ colind, const uint64_t *__restrict__ vals,
                   const uint64_t *__restrict__ x, uint64_t *__restrict__ y, const uint32_t *__restrict__ seg_row,
                   uint32_t rows, const uint32_t *__restrict__ warp_seg, uint32_t *__restrict__ carry_row,
                   uint64_t *__restrict__ carry_val, uint32_t nwt, uint32_t chunk, uint32_t nchunks, uint32_t pf_dist)
{
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const uint64_t pol_keep = policy_evict_last(), pol_stream = policy_evict_first();
    const uint32_t nwarps = gridDim.x * (SK_THREADS / 32);
    __shared__ uint64_t s_close[SK_THREADS / 32][WK_ENTRIES];  // per warp: sums of the rows the current step closes
    uint64_t *sv = s_close[threadIdx.x >> 5];
#define B2_STREAM_PREFETCH()                                                                                              \
    if (pf_dist != 0 && lane == 0 && t + pf_dist < t1) {                                                                  \
        const uint64_t e = (uint64_t)(t + pf_dist) * WK_ENTRIES;                                                          \
        asm volatile("cp.async.bulk.prefetch.L2.global.L2::cache_hint [%0], %1, %2;" ::"l"(vals + e), "r"(WK_ENTRIES * 8u), "l"(pol_stream) : "memory");   \
        asm volatile("cp.async.bulk.prefetch.L2.global.L2::cache_hint [%0], %1, %2;" ::"l"(colind + e), "r"(WK_ENTRIES * 4u), "l"(pol_stream) : "memory"); \
    }
#define B2_STREAM_STEP(C, P) \
    stream_step<DT, PERSIST, READ_Y, HAS_EMPTY>(C, P, x, y, seg_row, rows, lane, lt, pol_keep, sv, segs_base, wcarry, row_n, y_n)
    for (uint32_t vc = blockIdx.x * (SK_THREADS / 32) + (threadIdx.x >> 5); vc < nchunks; vc += nwarps) {
        uint32_t t = vc * chunk;
        const uint32_t t1 = min(t + chunk, nwt);
        uint32_t segs_base = __ldg(warp_seg + t);
        uint64_t wcarry = 0;
        // pipeline prologue: row id and old y of slot `lane` of the first step
        const uint32_t first_seg = segs_base + (uint32_t)lane - 1u;
        uint32_t row_n = first_seg != 0xFFFFFFFFu ? segment_row<HAS_EMPTY>(seg_row, rows, first_seg) : NO_ROW;
        uint64_t y_n = (READ_Y && row_n < rows) ? y[row_n] : 0ull;
        const uint32_t *cp = colind + (uint64_t)t * WK_ENTRIES + lane * 8;
        const uint64_t *vp = vals + (uint64_t)t * WK_ENTRIES + lane * 8;
        uint32_t ca[8];
        uint64_t pa[8];
        stream_load(ca, pa, cp, vp);
        if (DB) {
            uint32_t cb[8];
            uint64_t pb[8];
            while (true) {
                if (t + 1 < t1) stream_load(cb, pb, cp + WK_ENTRIES, vp + WK_ENTRIES);
                B2_STREAM_PREFETCH();
                B2_STREAM_STEP(ca, pa);
                if (++t >= t1) break;
                cp += WK_ENTRIES; vp += WK_ENTRIES;
                if (t + 1 < t1) stream_load(ca, pa, cp + WK_ENTRIES, vp + WK_ENTRIES);
                B2_STREAM_PREFETCH();
                B2_STREAM_STEP(cb, pb);
                if (++t >= t1) break;
                cp += WK_ENTRIES; vp += WK_ENTRIES;
            }
        } else {
            while (true) {
                B2_STREAM_PREFETCH();
                B2_STREAM_STEP(ca, pa);
                if (++t >= t1) break;
                cp += WK_ENTRIES; vp += WK_ENTRIES;
                stream_load(ca, pa, cp, vp);
            }
        }
#undef B2_STREAM_STEP
#undef B2_STREAM_PREFETCH
        if (lane == 0) {  // carry-out: the row left open at the end of the chunk
            uint32_t row = NO_ROW;
            if (B200SPMV_ABLATE == 0 && segs_base != 0) {
                row = HAS_EMPTY ? __ldg(seg_row + segs_base - 1) : segs_base - 1;
                if (row >= rows) row = NO_ROW;
            }
            carry_row[vc] = row;
            carry_val[vc] = wcarry;
        }
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

// Flagged format for kernels 3 and 6 (needs bit 31 of the column ids free; otherwise only the row-sequential
// kernel runs).
// ---- b200spmv_exec_noalloc: carry scratch that is never allocated ---------------------------------
// The tile carries of kernel 3 (one row id + one partial sum per 2048-entry tile) normally live in arrays that
// setup cudaMalloc's.  For callers that must not allocate, the same kernel can write them into this file-scope
// array instead: NOALLOC_MAX_TILES carry values (8 B each) followed by as many carry rows (4 B each) = 3 MB that
// every GPU which loads the library holds until the process exits.  In HOST code the name below is only a
// stand-in for the symbol, not a device address; the address is asked for once per device at setup
// (noalloc_scratch_address, called from b2_build_flags with the context's device current).
constexpr uint32_t NOALLOC_MAX_TILES = 1u << 18;  // per column slab: 2^18 tiles x 2048 = 2^29 stored entries
static __device__ __align__(256) unsigned char g_noalloc_scratch[(size_t)NOALLOC_MAX_TILES * 12];
static void *g_noalloc_addr[64] = {nullptr};      // by device ordinal

static int noalloc_scratch_address(b200spmv_ctx *ctx)
{
    const int dev = ctx->device;
    void *p = (dev >= 0 && dev < 64) ? g_noalloc_addr[dev] : nullptr;
    if (!p) {
        B2_CHECK(ctx, cudaSetDevice(dev));
        B2_CHECK(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (dev >= 0 && dev < 64) g_noalloc_addr[dev] = p;
    }
    ctx->noalloc_scratch = p;
    return 0;
}

int b2_build_flags(b200spmv_ctx *ctx)
{
    DeviceCsr &m = ctx->csr;
    cudaStream_t s = ctx->stream;
    b2_free(m.seg_row); b2_free(m.tile_seg); b2_free(m.warp_seg); b2_free(m.carry_row);
    { uint64_t *cv = (uint64_t *)m.carry_val; b2_free(cv); m.carry_val = nullptr; }
    ctx->flagged = false;
    ctx->colmask = 0xFFFFFFFFu;
    ctx->total_tiles = 0;
    if (ctx->cols > 0x80000000u || ctx->local_rows == 0) return 0;
    const uint32_t rows = ctx->local_rows;
    const size_t stride = ctx->rowptr_stride;
    uint32_t total = 0;
    for (int sl = 0; sl < ctx->slabs; sl++) {
        ctx->h_ftile_base[sl] = total;
        total += (uint32_t)((ctx->h_slab_off[sl + 1] - ctx->h_slab_off[sl]) / B2_FLAG_TILE);
    }
    ctx->h_ftile_base[ctx->slabs] = total;
    ctx->total_tiles = total;
    B2_TRY(b2_alloc(ctx, &m.seg_row, stride * ctx->slabs));
    B2_TRY(b2_alloc(ctx, &m.tile_seg, (size_t)total + ctx->slabs));
    B2_TRY(b2_alloc(ctx, &m.warp_seg, (size_t)total * WK_PER_TILE + ctx->slabs));
    // carries: one slot per 256-entry warp step serves every kernel (kernel 3 uses one per tile, kernel 6 one per chunk)
    B2_TRY(b2_alloc(ctx, &m.carry_row, (size_t)total * WK_PER_TILE + 1));
    { uint64_t *cv = nullptr; B2_TRY(b2_alloc(ctx, &cv, (size_t)total * WK_PER_TILE + 1)); m.carry_val = cv; }
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
    B2_TRY(noalloc_scratch_address(ctx));
    ctx->flagged = true;
    ctx->colmask = 0x7FFFFFFFu;
    return 0;
}

// Does this slab run the thread-per-row kernel?  Only slabs without a long row piece qualify (dcsr_maxrow).  By
// default the answer was measured at setup (b2_tune_slab_kernels); option dcsr_x16 >= 0 replaces the measurement
// by a fixed rule (0 = never, n = fewer than n/16 entries per non-empty row).
static bool dcsr_candidate(const b200spmv_ctx *ctx, int sl)
{
    return ctx->h_nseg[sl] > 0 && ctx->h_maxrow[sl] <= (uint32_t)ctx->opt_dcsr_maxrow;
}
static bool dcsr_slab(const b200spmv_ctx *ctx, int sl)
{
    if (ctx->tune_force) return ctx->tune_force == 2 && (ctx->tune_mask >> sl & 1);
    if (!dcsr_candidate(ctx, sl)) return false;
    if (ctx->opt_dcsr_x16 < 0) return ctx->h_slab_dcsr[sl] == 1;
    return ctx->opt_dcsr_x16 > 0 && (uint64_t)ctx->h_slab_nnz[sl] * 16 < (uint64_t)ctx->h_nseg[sl] * ctx->opt_dcsr_x16;
}
// Will b2_spmv_launch cut the last slab's kernel into row chunks (and record ev_rowchunk[])?  The flagged layout
// only: the tiles of kernel 3 are consecutive pieces of the entry stream, hence of the rows (whichever of kernels
// 3 and 6 is selected, the chunked last slab of b200spmv_exec runs kernel 3).
bool b2_last_slab_chunkable(const b200spmv_ctx *ctx, int flags)
{
    return ctx->chunks_ready && ctx->flagged && !(flags & B200SPMV_STRICT_ORDER) &&
           ctx->order_n == 0 && ctx->local_rows > 0 && !dcsr_slab(ctx, ctx->slabs - 1);
}

// Is the persisting window (and the evict_last hint on the gathers) on for this call?
static bool want_persist(const b200spmv_ctx *ctx, int flags)
{
    if (flags & B200SPMV_NO_PERSIST) return false;
    if (ctx->tune_persist >= 0) return ctx->tune_persist != 0;
    if (ctx->opt_persist >= 0) return ctx->opt_persist != 0;
    return ctx->persist_tuned;
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

// template dispatch over the three run-time booleans (PERSIST, READ_Y, HAS_EMPTY)
#define B2_DISPATCH3(p, r, e, CALL)                                      \
    do {                                                                 \
        if (p) {                                                         \
            if (r) { if (e) CALL(true, true, true); else CALL(true, true, false); }      \
            else { if (e) CALL(true, false, true); else CALL(true, false, false); }      \
        } else {                                                         \
            if (r) { if (e) CALL(false, true, true); else CALL(false, true, false); }    \
            else { if (e) CALL(false, false, true); else CALL(false, false, false); }    \
        }                                                                \
    } while (0)

static bool prof_begin(b200spmv_ctx *ctx, cudaStream_t s)
{
    if (!ctx->opt_profile || ctx->prof_n + 2 > 128) return false;
    if (!ctx->prof_ev[ctx->prof_n]) { cudaEventCreate(&ctx->prof_ev[ctx->prof_n]); cudaEventCreate(&ctx->prof_ev[ctx->prof_n + 1]); }
    cudaEventRecord(ctx->prof_ev[ctx->prof_n], s);
    return true;
}
static void prof_end(b200spmv_ctx *ctx, cudaStream_t s)
{
    cudaEventRecord(ctx->prof_ev[ctx->prof_n + 1], s);
    ctx->prof_n += 2;
}

template <int DT>
static int launch_dtype(b200spmv_ctx *ctx, int flags, cudaStream_t s)
{
    const DeviceCsr &m = ctx->csr;
    const bool strict = (flags & B200SPMV_STRICT_ORDER) || !ctx->flagged;
    const bool persist = want_persist(ctx, flags);
    const bool accumulate = flags & B200SPMV_ACCUMULATE;
    const uint64_t *xfull = (const uint64_t *)ctx->d_x;
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
                x_prefetch_kernel<<<(unsigned)((nthr + 127) / 128), 128, 0, s>>>((const char *)xfull + b0, b1 - b0);
                ctx->launches += 1;
            }
        }
        if (ctx->wait_slab_events) B2_CHECK(ctx, cudaStreamWaitEvent(s, ctx->ev_slab[sl], 0));
        if (strict) {
            const unsigned grid = b2_div_up(rows, 256);
            if (read_y) spmv_rowseq_kernel<DT, true><<<grid, 256, 0, s>>>(rowptr, m.colind, vals, xfull, y, rows, ctx->colmask);
            else spmv_rowseq_kernel<DT, false><<<grid, 256, 0, s>>>(rowptr, m.colind, vals, xfull, y, rows, ctx->colmask);
            ctx->launches += 1;
            B2_CHECK(ctx, cudaGetLastError());
            continue;
        }
        const uint32_t tb = ctx->h_ftile_base[sl];
        const uint32_t nt = ctx->h_ftile_base[sl + 1] - tb;
        const bool has_empty = ctx->h_nseg[sl] < rows;
        if (!read_y && has_empty) B2_CHECK(ctx, cudaMemsetAsync(y, 0, (size_t)rows * 8, s));  // rows no flag will touch
        // packed exchange (multi.cu): a slab of another rank's block reads the packed block through the renumbered indices
        const bool packed = ctx->packed_active && ordered && (sl < ctx->own_lo || sl >= ctx->own_hi);
        const uint32_t *colbase = packed ? ctx->colind_packed : m.colind;
        const uint64_t *x = packed ? ctx->d_unpack : xfull;
        const uint32_t *ci = colbase + ctx->h_slab_off[sl];
        const uint64_t *va = vals + ctx->h_slab_off[sl];
        const uint32_t *segrow = m.seg_row + sl * stride, *tseg = m.tile_seg + tb + sl;
        const bool prof = prof_begin(ctx, s);
        // hypersparse slab (fewer than opt_dcsr_x16/16 entries per non-empty row and no long row: on a skewed
        // matrix one thread with a few thousand entries holds the whole launch up): thread per non-empty row
        const uint32_t nseg = ctx->h_nseg[sl];
        if (dcsr_slab(ctx, sl)) {
            const unsigned grid = b2_div_up(nseg, 256);
            if (persist) {
                if (read_y) spmv_dcsr_kernel<DT, true, true><<<grid, 256, 0, s>>>(rowptr, colbase, vals, x, y, segrow, nseg, ctx->colmask);
                else spmv_dcsr_kernel<DT, true, false><<<grid, 256, 0, s>>>(rowptr, colbase, vals, x, y, segrow, nseg, ctx->colmask);
            } else {
                if (read_y) spmv_dcsr_kernel<DT, false, true><<<grid, 256, 0, s>>>(rowptr, colbase, vals, x, y, segrow, nseg, ctx->colmask);
                else spmv_dcsr_kernel<DT, false, false><<<grid, 256, 0, s>>>(rowptr, colbase, vals, x, y, segrow, nseg, ctx->colmask);
            }
            if (prof) prof_end(ctx, s);
            ctx->launches += 1;
            B2_CHECK(ctx, cudaGetLastError());
            continue;
        }
        // For b200spmv_exec the last slab goes out in Y_CHUNKS pieces of consecutive tiles (kernel 3), each followed
        // by its own fix-up and an event: the rows a piece completes can leave for the host while the pieces after it
        // still run (common.cuh, h_chunk_row).
        const bool chunked = ctx->want_row_chunks && pos == ctx->slabs - 1 && b2_last_slab_chunkable(ctx, flags);
        if (ctx->opt_kernel == 6 && !chunked) {
            const uint32_t nwt = nt * WK_PER_TILE;
            const uint32_t *wseg = m.warp_seg + (size_t)tb * WK_PER_TILE + sl;
            uint32_t *wcrow = m.carry_row + (size_t)tb * WK_PER_TILE;
            uint64_t *wcval = (uint64_t *)m.carry_val + (size_t)tb * WK_PER_TILE;
            const bool db = ctx->opt_stream_db != 0;
            const uint32_t occ = ctx->opt_stream_occ > 0 ? (uint32_t)ctx->opt_stream_occ : (db ? 2u : 3u);
            const uint32_t warps = (uint32_t)ctx->sms * occ * (SK_THREADS / 32);
            const uint32_t chunk = std::max<uint32_t>(1u, b2_div_up(nwt, (uint64_t)warps * (uint32_t)std::max(1, ctx->opt_stream_vchunks)));
            const uint32_t nchunks = b2_div_up(nwt, chunk);
            const uint32_t grid = std::min<uint32_t>((uint32_t)ctx->sms * occ, b2_div_up(nchunks, SK_THREADS / 32));
            const uint32_t pf = (uint32_t)ctx->opt_stream_pf;
#define B2_CALL_STREAM(P, R, E)                                                                                            \
    do {                                                                                                                   \
        if (db) spmv_stream_kernel<DT, P, R, E, true><<<grid, SK_THREADS, 0, s>>>(ci, va, x, y, segrow, rows, wseg, wcrow, wcval, nwt, chunk, nchunks, pf); \
        else spmv_stream_kernel<DT, P, R, E, false><<<grid, SK_THREADS, 0, s>>>(ci, va, x, y, segrow, rows, wseg, wcrow, wcval, nwt, chunk, nchunks, pf);  \
    } while (0)
            B2_DISPATCH3(persist, read_y, has_empty, B2_CALL_STREAM);
#undef B2_CALL_STREAM
            if (prof) prof_end(ctx, s);
            spmv_fixup_kernel<DT><<<b2_div_up(nchunks, 256), 256, 0, s>>>(wcrow, wcval, nchunks, y);
            ctx->launches += 2;
            B2_CHECK(ctx, cudaGetLastError());
            continue;
        }
        uint32_t *crow = m.carry_row + tb;
        uint64_t *cval = (uint64_t *)m.carry_val + tb;
        const int pieces = chunked ? Y_CHUNKS : 1;
        for (int k = 0; k < pieces; k++) {
            const uint32_t t0 = chunked ? ctx->h_chunk_tile[k] : 0, t1 = chunked ? ctx->h_chunk_tile[k + 1] : nt;
            const uint32_t pn = t1 - t0;
            if (pn > 0) {
                const uint32_t *pci = ci + (size_t)t0 * B2_FLAG_TILE, *ptseg = tseg + t0;
                const uint64_t *pva = va + (size_t)t0 * B2_FLAG_TILE;
                uint32_t *pcrow = crow + t0;
                uint64_t *pcval = cval + t0;
#define B2_CALL_FLAG(P, R, E) \
    spmv_flag_kernel<DT, P, R, E, 8><<<pn, B2_FLAG_TILE / 8, 0, s>>>(pci, pva, x, y, segrow, rows, ptseg, pcrow, pcval, (uint32_t)ctx->opt_prefetch_tiles)
                B2_DISPATCH3(persist, read_y, has_empty, B2_CALL_FLAG);
#undef B2_CALL_FLAG
                ctx->launches += 1;
            }
            if (k == pieces - 1 && prof) prof_end(ctx, s);
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
        B2_CHECK(ctx, cudaGetLastError());
    }
    ctx->last_mode = strict ? 0 : ctx->opt_kernel;
    return 0;
}

// Flagged kernel or thread-per-row kernel?  Which one is faster on a slab of short rows depends on more than the
// mean row length: uniform 4 per row favours the thread-per-row kernel by 8 %, the three slabs of banded(11)+5
// (5.85 per non-empty row, but a mixture of 12.7 inside the diagonal block and 1.7 outside) favour the flagged kernel by
// 50 %, and with 8 slabs of the same matrix it is the other way round by 43 % (profiles/r02_sweep_dcsr.txt).  So
// when some slab qualifies (no long row piece, fewer than 8 entries per non-empty row, big enough to time), setup ends
// with whole steps y += A*x timed slab by slab (the per-launch events of option "profile"): three with the flagged
// kernel everywhere, three with the thread-per-row kernel on the qualifying slabs, the first of each untimed.  Whole
// steps, because a slab run on its own over and over finds its slice of x already in L2, which it never does in
// use.  The thread-per-row kernel is kept where it wins by more than 5 % (a tie must not flip between two setups of
// the same matrix).  x and y are zero at this point; y is zeroed again afterwards (inf times 0 would leave a NaN).
int b2_tune_slab_kernels(b200spmv_ctx *ctx)
{
    memset(ctx->h_slab_dcsr, 0, sizeof ctx->h_slab_dcsr);
    ctx->tunetime_ns = 0;
    ctx->tune_mask = 0;
    ctx->persist_tuned = true;
    if (!ctx->flagged || ctx->local_rows == 0) return 0;
    for (int sl = 0; sl < ctx->slabs; sl++) {
        const uint64_t nnz = ctx->h_slab_nnz[sl], nseg = ctx->h_nseg[sl];
        if (!dcsr_candidate(ctx, sl) || nnz >= 8 * nseg) continue;
        if (nnz < (1u << 18)) ctx->h_slab_dcsr[sl] = nnz < 6 * nseg;  // launch-latency sized: the fixed rule of round 1
        else ctx->tune_mask |= 1ull << sl;
    }
    cudaStream_t s = ctx->stream;
    const auto t_begin = std::chrono::steady_clock::now();
    // The access-policy window first: it keeps more of the gathered slice of x in L2 (profiles/r02_l2_ab.txt) but takes
    // the set-aside away from y, whose rows the kernels update by reductions in L2.  Measured both ways, three whole
    // steps at a time (the first untimed), twice each: band, R-MAT and one-slab matrices gain 3-6 % from the window, C3 and short-row
    // slabs lose 2-6 % (profiles/r02_y_red_ab.txt).  Kept on unless off is more than 1 % faster.
    ctx->persist_tuned = true;
    bool ran_persist = false;
    if (ctx->opt_persist < 0 && ctx->max_persist_bytes > 0 && (uint64_t)ctx->cols * 8 >= (8u << 20) && ctx->local_nnz >= (1u << 22)) {
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        B2_CHECK(ctx, cudaEventCreate(&e0));
        B2_CHECK(ctx, cudaEventCreate(&e1));
        float t[2] = {1e30f, 1e30f};
        int rc = 0;
        // OFF first: lines the window has marked persisting stay in the set-aside after it is switched off, so an
        // "off" measured after an "on" still runs with x pinned (on a rank's block of the 8-way split that hid a 9 %
        // difference); between the tries the persisting lines are dropped
        for (int round = 0; round < 4 && !rc; round++) {  // off, on, off, on: the best of two tries each
            const int on = round & 1;
            if (!on) { cudaStreamSynchronize(s); cudaCtxResetPersistingL2Cache(); }
            ctx->tune_persist = on;
            ctx->tune_force = 1;
            rc = b2_spmv_launch(ctx, B200SPMV_ACCUMULATE, s);
            cudaEventRecord(e0, s);
            for (int rep = 0; rep < 3 && !rc; rep++) rc = b2_spmv_launch(ctx, B200SPMV_ACCUMULATE, s);
            cudaEventRecord(e1, s);
            if (!rc && cudaEventSynchronize(e1) != cudaSuccess) rc = b2_fail(ctx, B200SPMV_ECUDA, "persisting-window measurement failed");
            float ms = 0.f;
            if (!rc && cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess) t[on] = std::min(t[on], ms);
        }
        ctx->tune_persist = -1;
        ctx->tune_force = 0;
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        if (rc) return rc;
        ctx->persist_tuned = !(t[0] < 0.99f * t[1]);
        if (!ctx->persist_tuned) { cudaStreamSynchronize(s); cudaCtxResetPersistingL2Cache(); }
        ran_persist = true;
    }
    if (!ctx->tune_mask || ctx->opt_dcsr_x16 >= 0) {  // no slab to measure / a fixed rule was asked for
        if (ran_persist) {
            B2_CHECK(ctx, cudaMemsetAsync(ctx->d_y, 0, (size_t)ctx->rows * 8, s));
            B2_CHECK(ctx, cudaStreamSynchronize(s));
            ctx->launches = 0;
            ctx->tunetime_ns = (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t_begin).count();
        }
        return 0;
    }
    const int profile_was = ctx->opt_profile;
    uint64_t best[3][64];
    memset(best, 0xFF, sizeof best);
    int rc = 0;
    for (int which = 1; which <= 2 && !rc; which++) {
        ctx->tune_force = which;
        for (int rep = 0; rep < 3 && !rc; rep++) {
            ctx->opt_profile = rep > 0;
            ctx->prof_n = 0;
            memset(ctx->prof_slab_ns, 0, sizeof ctx->prof_slab_ns);
            rc = b2_spmv_launch(ctx, B200SPMV_ACCUMULATE, s);
            if (!rc && cudaStreamSynchronize(s) != cudaSuccess) rc = b2_fail(ctx, B200SPMV_ECUDA, "slab kernel measurement failed");
            if (rc || rep == 0) continue;
            b2_profile_collect(ctx);
            for (int sl = 0; sl < ctx->slabs; sl++) best[which][sl] = std::min(best[which][sl], ctx->prof_slab_ns[sl]);
        }
    }
    ctx->tune_force = 0;
    ctx->opt_profile = profile_was;
    ctx->prof_n = 0;
    ctx->merge_kernel_ns = 0;
    ctx->merge_kernel_launches = 0;
    memset(ctx->prof_slab_ns, 0, sizeof ctx->prof_slab_ns);
    if (rc) return rc;
    for (int sl = 0; sl < ctx->slabs; sl++)
        if (ctx->tune_mask >> sl & 1) ctx->h_slab_dcsr[sl] = (double)best[2][sl] < 0.95 * (double)best[1][sl];
    B2_CHECK(ctx, cudaMemsetAsync(ctx->d_y, 0, (size_t)ctx->rows * 8, s));
    B2_CHECK(ctx, cudaStreamSynchronize(s));
    ctx->launches = 0;
    ctx->tunetime_ns = (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t_begin).count();
    return 0;
}

// Fold the recorded kernel event pairs into merge_kernel_ns (call after a sync).
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
    const bool persist = want_persist(ctx, flags);
    if (persist != ctx->persist_set) {
        B2_TRY(set_persist_window(ctx, s, ctx->d_x, (size_t)ctx->cols * 8, persist));
        // switching the window off must also give the set-aside back (lines marked persisting outlive the window)
        if (!persist && ctx->persist_set) { cudaStreamSynchronize(s); cudaCtxResetPersistingL2Cache(); }
        ctx->persist_set = persist;
    }
    return 0;
}

int b2_spmv_launch(b200spmv_ctx *ctx, int flags, cudaStream_t s)
{
    B2_TRY(b2_spmv_prepare(ctx, flags, s));
    return ctx->dtype == B200SPMV_F64 ? launch_dtype<B200SPMV_F64>(ctx, flags, s)
                                      : launch_dtype<B200SPMV_U64>(ctx, flags, s);
}

// ---- the no-allocation entry point (no counterpart in the reference) --
// y = A*x or y += A*x on DEVICE vectors the caller passes, with kernel 3 on every column slab and the tile carries in
// the static scratch above: no cudaMalloc, no cudaFree, no event or stream creation, no stream attribute -- only
// kernel launches (and one memset of y for y = A*x on a matrix with empty rows) on the given stream.  The matrix is
// the one the context was set up with.  The kernel writes every tile's carry before the fix-up reads it, so nothing
// in the scratch has to start at zero; it keeps the last call's carries.  All calls on one GPU share the scratch:
// two calls whose kernels overlap on the GPU (different streams, different contexts) race.
template <int DT>
static int launch_noalloc(b200spmv_ctx *ctx, const uint64_t *x, uint64_t *yfull, int flags, cudaStream_t s)
{
    const DeviceCsr &m = ctx->csr;
    const uint32_t rows = ctx->local_rows;
    if (rows == 0) return 0;
    uint64_t *y = yfull + ctx->row_begin;
    uint64_t *cval = (uint64_t *)ctx->noalloc_scratch;
    uint32_t *crow = (uint32_t *)((unsigned char *)ctx->noalloc_scratch + (size_t)NOALLOC_MAX_TILES * 8);
    const size_t stride = ctx->rowptr_stride;
    for (int sl = 0; sl < ctx->slabs; sl++) {
        const uint32_t tb = ctx->h_ftile_base[sl], nt = ctx->h_ftile_base[sl + 1] - tb;
        const bool read_y = (flags & B200SPMV_ACCUMULATE) || sl > 0;
        const bool has_empty = ctx->h_nseg[sl] < rows;
        if (!read_y && has_empty) B2_CHECK(ctx, cudaMemsetAsync(y, 0, (size_t)rows * 8, s));  // rows no flag will touch
        if (nt == 0) continue;
        const uint32_t *ci = m.colind + ctx->h_slab_off[sl];
        const uint64_t *va = (const uint64_t *)m.vals + ctx->h_slab_off[sl];
        const uint32_t *segrow = m.seg_row + sl * stride, *tseg = m.tile_seg + tb + sl;
#define B2_CALL_NOALLOC(P, R, E) \
    spmv_flag_kernel<DT, P, R, E, 8><<<nt, B2_FLAG_TILE / 8, 0, s>>>(ci, va, x, y, segrow, rows, tseg, crow, cval, 0u)
        B2_DISPATCH3(false, read_y, has_empty, B2_CALL_NOALLOC);
#undef B2_CALL_NOALLOC
        spmv_fixup_kernel<DT><<<b2_div_up(nt, 256), 256, 0, s>>>(crow, cval, nt, y);
        B2_CHECK(ctx, cudaGetLastError());
    }
    return 0;
}

extern "C" int b200spmv_exec_noalloc(b200spmv_ctx *ctx, const void *x_dev, void *y_dev, int flags, void *stream)
{
    if (!ctx) return B200SPMV_EINVAL;
    if (!ctx->ready) return b2_fail(ctx, B200SPMV_ESTATE, "exec_noalloc before setup");
    if (!x_dev || !y_dev) return b2_fail(ctx, B200SPMV_EINVAL, "null vector");
    if (flags & ~B200SPMV_ACCUMULATE) return b2_fail(ctx, B200SPMV_EINVAL, "exec_noalloc takes B200SPMV_ACCUMULATE or 0");
    if (!ctx->flagged || !ctx->noalloc_scratch) return b2_fail(ctx, B200SPMV_ESTATE, "exec_noalloc needs the flagged layout");
    for (int sl = 0; sl < ctx->slabs; sl++) {
        const uint32_t nt = ctx->h_ftile_base[sl + 1] - ctx->h_ftile_base[sl];
        if (nt > NOALLOC_MAX_TILES) {
            char buf[200];
            snprintf(buf, sizeof buf, "exec_noalloc: column slab %d has %u tiles of 2048 entries, the static scratch holds %u (2^29 entries per slab)",
                     sl, nt, NOALLOC_MAX_TILES);
            ctx->err = buf;
            return B200SPMV_EINVAL;
        }
    }
    B2_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    const int rc = ctx->dtype == B200SPMV_F64 ? launch_noalloc<B200SPMV_F64>(ctx, (const uint64_t *)x_dev, (uint64_t *)y_dev, flags, s)
                                              : launch_noalloc<B200SPMV_U64>(ctx, (const uint64_t *)x_dev, (uint64_t *)y_dev, flags, s);
    if (rc) return rc;
    if (!stream) B2_CHECK(ctx, cudaStreamSynchronize(s));  // the caller's own stream is the caller's to wait on
    return 0;
}
