"""Small solves that reach every kernel variant, for compute-sanitizer (tools/gpu.sh sanitize)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import auctionalgorithm_b200 as ab
from oracle import pyoracle as po
import parity

h = ab.Handle(0)
which = sys.argv[1:] or ["dense", "pairs", "sparse", "batch", "fused"]
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
    # two-lane pass, dynamic hand-out over more than one reservation, takeover by cap (forced at 8 bids) and by a
    # real price war (problem 1117884 of the bench shards: 1483 bids), half last chunks (cols % 8 == 4)
    a = np.stack([po.generate(32, 48, 1, problem=b) for b in [1117884] + list(range(200))])
    ref, failed = po.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    for cap in (0, 8, -1):
        assert np.array_equal(h.solve_batch(a, batch_bid_cap=cap), ref)
    b = np.stack([po.generate(20, 44, 2, problem=k) for k in range(100)])
    ref, _ = po.oracle_batch(b, use_double=True, n_threads=4)
    assert np.array_equal(h.solve_batch(b, batch_bid_cap=4), ref)
    c = np.stack([po.generate(9, 13, 3, problem=k) for k in range(50)])     # streaming kernel (cols % 4 != 0)
    ref, _ = po.oracle_batch(c, use_double=True, n_threads=4)
    assert np.array_equal(h.solve_batch(c), ref)
    print("batch", a.shape, b.shape, c.shape, flush=True)
if "fused" in which:
    # column-sharded solve, three ranks on this GPU: peer stores into the mailboxes, release / acquire flags
    import torch
    from auctionalgorithm_b200 import sharded
    hs = [ab.Handle(0) for _ in range(3)]
    sharded.connect_fused_local(hs, 256)
    for rows, cols in ((200, 300), (256, 256)):
        a = po.generate(rows, cols, 1)
        parts = sharded.split_columns(cols, 3)
        shards = [torch.from_numpy(np.ascontiguousarray(a[:, c0:c0 + cl])).cuda() for c0, cl in parts]
        res = sharded.solve_fused_local(hs, shards, rows, parts, [cl for _, cl in parts], cols)
        assert parity.is_valid_assignment(res[0][0].cpu().numpy(), cols)
        print("fused", rows, cols, res[0][2]["rounds"], flush=True)
    for x in hs:
        x.close()
print("done")
