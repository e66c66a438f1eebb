"""Warm start (options.prices_in), price export and the torch ops, on the GPU."""
import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu


def _oracle_r2c(oracle_lib, a):
    out = oracle_lib.oracle_solve(a.astype(np.float64))
    r2c = np.full(a.shape[0], -1, np.int64)
    r2c[out["x"]] = out["y"]
    return r2c


@pytest.mark.parametrize("rows,cols", [(512, 512), (300, 300), (200, 300), (1024, 1536)])
def test_warm_start_equals_cold_solve(handle, oracle_lib, rows, cols):
    # frame 1: cold solve, prices exported; frame 2: a few percent of the coefficients move a little (a tracker's next
    # frame); the warm solve starts from frame 1's prices and must return what the cold solve of frame 2 returns
    rng = np.random.default_rng(3)
    a1 = oracle_lib.generate(rows, cols, 5)
    r1, p1 = handle.solve(a1, want_prices=True)
    a2 = a1.copy()
    idx = rng.random(a1.shape) < 0.02
    a2[idx] = np.clip(a2[idx] + rng.normal(0, 0.01, idx.sum()).astype(np.float32), 2.0 ** -24, 1.0).astype(np.float32)
    cold, pc = handle.solve(a2, want_prices=True)
    cold_rounds = handle.last_stats["rounds"]
    warm, pw = handle.solve(a2, want_prices=True, prices_in=p1)
    st = handle.last_stats
    assert np.array_equal(warm, cold)
    assert parity.check_parity(a2, warm, pw, _oracle_r2c(oracle_lib, a2), 1e-5 / cols) == "exact"
    if rows == cols:
        assert st["warm_fallback"] == 0
        assert st["rounds"] <= cold_rounds * 1.1, (st["rounds"], cold_rounds)   # measured: 10-30 % fewer
    # rows < cols: a column that frame 1 priced and frame 2 leaves unassigned breaks the post-condition of
    # AuctionCommon.h:347-354; the floor-raise repair (stats.warm_fallback counts its iterations) restores it
    # the same matrix again from its own final prices, one phase at the final eps: hardly anything left to bid for
    again, _ = handle.solve(a2, want_prices=True, prices_in=pw, eps_scaling=0)
    assert np.array_equal(again, cold)
    if rows == cols:
        assert handle.last_stats["warm_fallback"] == 0
        assert handle.last_stats["rounds"] <= max(40, cold_rounds // 10), handle.last_stats["rounds"]


def test_warm_start_with_hostile_prices_on_an_asymmetric_problem_is_repaired(handle, oracle_lib):
    # prices that belong to nothing: the forward auction still ends with eps-CS, but unassigned columns may stay
    # priced above assigned ones (AuctionCommon.h:347-354 would not hold): detected and repaired by the floor raise
    rows, cols = 200, 300
    a = oracle_lib.generate(rows, cols, 2)
    g = np.load(f"{parity.GOLDEN_DIR}/auction_200x300_s2.npz")
    bad = np.random.default_rng(1).random(cols) * 0.9
    r2c, prices = handle.solve(a, want_prices=True, prices_in=bad)
    assert handle.last_stats["warm_fallback"] >= 1
    assert parity.check_parity(a, r2c, prices, parity.golden_row_to_col(g), 1e-5 / cols) == "exact"
    import auctionalgorithm_b200 as ab
    for broken in (np.full(cols, np.nan), -np.ones(cols)):
        with pytest.raises(ab.B200LsapError) as e:
            handle.solve(a, prices_in=broken)
        assert e.value.status == ab.ERR_ARG
    assert np.array_equal(handle.solve(a), parity.golden_row_to_col(g))   # the handle is still good


def test_torch_ops(handle, oracle_lib):
    import torch
    import auctionalgorithm_b200.torch_op  # noqa: F401  registers torch.ops.b200lsap.*
    g = np.load(parity.GOLDEN_DIR + "/batch_256x32x48_s1.npz")
    a = np.stack([oracle_lib.generate(32, 48, 1, problem=b) for b in range(256)])
    t = torch.from_numpy(a).cuda()
    out = torch.ops.b200lsap.solve(t)
    assert out.dtype == torch.int32 and out.shape == (256, 32) and out.is_cuda
    assert np.array_equal(out.cpu().numpy(), g["row_to_col"])
    # on a side stream: the op follows torch's current stream
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        out2 = torch.ops.b200lsap.solve(t * 1.0)
    s.synchronize()
    assert torch.equal(out, out2)
    # problems too large for the batched kernel go through the dense solver one by one
    big = np.stack([oracle_lib.generate(100, 130, 7, problem=b) for b in range(3)])
    got = torch.ops.b200lsap.solve(torch.from_numpy(big).cuda()).cpu().numpy()
    for b in range(3):
        assert np.array_equal(got[b].astype(np.int64), _oracle_r2c(oracle_lib, big[b]))
    # one dense problem, prices out, then back in
    g2 = np.load(parity.GOLDEN_DIR + "/auction_256x256_s2.npz")
    m = torch.from_numpy(oracle_lib.generate(256, 256, 2)).cuda()
    cols, prices = torch.ops.b200lsap.solve_dense(m, torch.empty(0, dtype=torch.float64, device="cuda"))
    assert np.array_equal(cols.cpu().numpy().astype(np.int64), parity.golden_row_to_col(g2))
    cols2, prices2 = torch.ops.b200lsap.solve_dense(m, prices)
    assert torch.equal(cols, cols2)
    with pytest.raises(RuntimeError):
        torch.ops.b200lsap.solve(t.cpu())
    with pytest.raises(RuntimeError):
        torch.ops.b200lsap.solve(t.double())
