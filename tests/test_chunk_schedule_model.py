"""Host-side model of the download pipeline's schedule (csrc/spmv.cu: kernel 3 launched in pieces of tiles,
spmv_fixup_range_kernel, chunk_rows_kernel; csrc/api.cu: b200spmv_exec).

The flagged kernel stores a row when it meets the NEXT row start, and leaves the partial sum of the row still open
at the end of a tile in carry_row/carry_val; a fix-up adds each run of carries to y.  Cut into pieces, the fix-up
of a carry must wait for the tile that stores its row, and the rows declared final after piece k must really be
final.  This restates those rules in plain Python on small random matrices (integers, so sums are exact) and
checks, for every piece boundary, what the device tests can only check at the end of the step."""
import numpy as np
import pytest

NO_ROW = 0xFFFFFFFF


def build(rng, rows, max_len, tile, empty_frac):
    """CSR row lengths -> padded entry stream with row-start flags, seg_row, tile_seg (b2_build_flags)."""
    lens = rng.integers(1, max_len + 1, size=rows)
    lens[rng.random(rows) < empty_frac] = 0
    rowptr = np.concatenate([[0], np.cumsum(lens)])
    nnz = int(rowptr[-1])
    padded = (nnz + tile) // tile * tile  # at least one pad entry: the sentinel that closes the last row
    vals = np.zeros(padded, dtype=np.int64)
    vals[:nnz] = rng.integers(-50, 50, size=nnz)
    flag = np.zeros(padded, dtype=bool)
    seg_row = []
    for r in range(rows):
        if lens[r]:
            flag[rowptr[r]] = True
            seg_row.append(r)
    flag[nnz] = True
    seg_row.append(NO_ROW)
    ntiles = padded // tile
    tile_seg = [int(flag[:t * tile].sum()) for t in range(ntiles + 1)]  # flags strictly before the tile's first entry
    truth = np.array([vals[rowptr[r]:rowptr[r + 1]].sum() for r in range(rows)], dtype=np.int64)
    return vals, flag, seg_row, tile_seg, ntiles, truth


def run_tile(t, tile, vals, flag, seg_row, tile_seg, rows, y, read_y, carry_row, carry_val):
    """spmv_flag_kernel on one tile: a flag closes the row open before it; the open tail goes to the carry."""
    seg = tile_seg[t]  # row starts before this tile: the open row on entry is segment seg - 1
    acc = 0
    for p in range(t * tile, (t + 1) * tile):
        if flag[p]:
            if seg > 0:
                row = seg_row[seg - 1]
                if row != NO_ROW and row < rows:
                    y[row] = (y[row] if read_y else 0) + acc
            seg += 1
            acc = 0
        acc += int(vals[p])
    row = seg_row[seg - 1] if seg > 0 else NO_ROW
    carry_row[t] = row if (row != NO_ROW and row < rows) else NO_ROW
    carry_val[t] = acc


def fixup_range(carry_row, carry_val, lo, hi, ntiles, y):
    """spmv_fixup_range_kernel: the thread at the END of a run of equal carry rows adds the run."""
    for e in range(lo, hi):
        row = carry_row[e]
        if row == NO_ROW:
            continue
        if e + 1 < ntiles and carry_row[e + 1] == row:
            continue
        h = e
        while h > 0 and carry_row[h - 1] == row:
            h -= 1
        y[row] += sum(carry_val[h:e + 1])


def chunk_rows(tile_seg, seg_row, ntiles, rows, chunks):
    """chunk_rows_kernel: piece boundaries in tiles and the row below which everything is final after each piece."""
    T, R = [], []
    for k in range(chunks + 1):
        t = ntiles * k // chunks
        r = 0
        if k == chunks:
            r = rows
        elif t > 0:
            c = tile_seg[t]
            r = seg_row[c - 1] if c else 0
            r = min(r, rows)
        T.append(t)
        R.append(r)
    return T, R


@pytest.mark.parametrize("read_y", [False, True])
@pytest.mark.parametrize("rows,max_len,tile,chunks,empty_frac", [
    (200, 3, 8, 4, 0.0), (200, 3, 8, 16, 0.3), (64, 40, 8, 8, 0.1), (5, 200, 8, 16, 0.0), (300, 1, 8, 16, 0.5),
    (37, 9, 16, 3, 0.2), (3, 2, 8, 16, 0.0), (400, 8, 8, 16, 0.0)])
def test_pieces_complete_the_rows_they_promise(rows, max_len, tile, chunks, empty_frac, read_y):
    rng = np.random.default_rng(rows * 1000 + max_len * 10 + chunks)
    for trial in range(20):
        vals, flag, seg_row, tile_seg, ntiles, truth = build(rng, rows, max_len, tile, empty_frac)
        y0 = rng.integers(-5, 5, size=rows).astype(np.int64) if read_y else np.zeros(rows, dtype=np.int64)
        want = truth + y0
        # a slab that overwrites clears the rows no flag will touch first (cudaMemsetAsync in b2_spmv_launch);
        # rows with entries hold stale data until they are stored
        y = y0.copy() if read_y else np.zeros(rows, dtype=np.int64)
        if not read_y:
            stale = rng.integers(100, 200, size=rows)
            nonempty = np.array([r for r in seg_row if r != NO_ROW], dtype=np.int64)
            y[nonempty] = stale[nonempty]
        carry_row, carry_val = [NO_ROW] * ntiles, [0] * ntiles
        T, R = chunk_rows(tile_seg, seg_row, ntiles, rows, chunks)
        assert R[0] == 0 and R[-1] == rows and T[-1] == ntiles and all(a <= b for a, b in zip(R, R[1:]))
        for k in range(chunks):
            t0, t1 = T[k], T[k + 1]
            for t in range(t0, t1):
                run_tile(t, tile, vals, flag, seg_row, tile_seg, rows, y, read_y, carry_row, carry_val)
            lo = 0 if k == 0 else max(t0, 1) - 1
            hi = ntiles if k == chunks - 1 else max(t1, 1) - 1
            if hi > lo:
                fixup_range(carry_row, carry_val, lo, hi, ntiles, y)
            # what b200spmv_exec downloads after piece k must be final now
            assert (y[:R[k + 1]] == want[:R[k + 1]]).all(), (trial, k)
        assert (y == want).all()


def test_fixing_up_a_piece_at_once_is_what_went_wrong():
    """The first cut of the pipeline ran the ordinary fix-up over a piece's own tiles: the carry of the piece's last
    tile was added before the next piece stored (overwrote) that row.  Keep the counter-example."""
    tile, rows = 8, 4
    # rows of 8 entries each: every tile boundary is a row boundary, every row is stored by the NEXT tile
    vals = np.ones(5 * tile, dtype=np.int64)
    vals[4 * tile:] = 0
    flag = np.zeros(5 * tile, dtype=bool)
    flag[::tile] = True
    seg_row = [0, 1, 2, 3, NO_ROW]
    tile_seg = [0, 1, 2, 3, 4, 5]
    y = np.full(rows, 99, dtype=np.int64)
    carry_row, carry_val = [NO_ROW] * 5, [0] * 5
    for t in (0, 1):
        run_tile(t, tile, vals, flag, seg_row, tile_seg, rows, y, False, carry_row, carry_val)
    for t in (0, 1):  # the wrong fix-up: runs headed inside the piece, added right away
        if carry_row[t] != NO_ROW:
            y[carry_row[t]] += carry_val[t]
    for t in (2, 3, 4):
        run_tile(t, tile, vals, flag, seg_row, tile_seg, rows, y, False, carry_row, carry_val)
    assert y[1] == 0  # row 1's carry (8) was added to stale data, then the store of tile 2 wiped it
