"""Small solves that reach every kernel variant, for compute-sanitizer (tools/gpu.sh sanitize)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import auctionalgorithm_b200 as ab
from oracle import pyoracle as po
import parity

h = ab.Handle(0)
which = sys.argv[1:] or ["dense", "pairs", "sparse", "batch"]
if "dense" in which:
    for rows, cols in ((200, 300), (256, 256), (700, 700)):
        a = po.generate(rows, cols, 1)
        r = h.solve(a)
        assert parity.is_valid_assignment(r, cols)
        r64 = h.solve(a.astype(np.float64), eps_start=h.last_stats["eps_first"])
        assert np.array_equal(r, r64)
        print("dense", rows, cols, h.last_stats["rounds"], flush=True)
if "pairs" in which:  # more rows than warps and a badly filled last wave: half rows per warp pair
    a = po.generate(5000, 5120, 2)
    r = h.solve(a)
    assert parity.is_valid_assignment(r, 5120)
    print("pairs 5000x5120", h.last_stats["rounds"], flush=True)
if "sparse" in which:
    for rows, cols, mod, thr in ((64, 64, 4, 1), (300, 300, 16, 13), (2048, 2048, 64, 60), (40, 40, 1, 0)):
        a, v, rp, ci = parity.make_csr(rows, cols, 3, np.float32, mod, thr)
        r = h.solve_csr(v, rp, ci, rows, cols)
        assert parity.is_valid_assignment(r, cols)
        print("sparse", rows, cols, len(v), h.last_stats["rounds"], flush=True)
if "batch" in which:
    a = np.stack([po.generate(32, 48, 1, problem=b) for b in range(64)])
    r = h.solve_batch(a)
    print("batch", r.shape, flush=True)
print("done")
