import sys, numpy as np
sys.path.insert(0, "spmv-vector-cache_b200"); sys.path.insert(0, "tests")
import b200spmv as b2
for rows, slabs, ctas in ((1 << 16, 1, 4), (1 << 20, 1, 4), (1 << 20, 1, 1), (1 << 22, 1, 4), (1 << 22, 3, 4)):
    e = b2.Engine(0)
    e.set_option("slabs", slabs); e.set_option("ctas_per_sm", ctas)
    e.setup_synthetic(0, rows, rows, 32, 0x5EED5EED, b2.U64)
    e.fill_x_synthetic(1)
    e.exec_device(1, b2.STRICT_ORDER); ys = e.get_y().copy()
    e.exec_device(1, 0); ym = e.get_y().copy()
    bad = np.nonzero(ys != ym)[0]
    print(rows, slabs, ctas, "tiles", rows * 32 // 2048, "bad rows", bad.size, bad[:10], (bad[:10] * 32) // 2048 if bad.size else "")
    e.close()
