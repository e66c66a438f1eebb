"""CPU tests: the C oracle against the reference-made golden vectors (tests/golden/, see
make_golden.py) and, when oracle/_ref exists, against the compiled reference itself."""
import numpy as np
import pytest

import parity


@pytest.mark.parametrize("path", parity.golden_cases(max_rows=2048))
def test_oracle_matches_golden_double(oracle_lib, path):
    g = np.load(path)
    rows, cols, seed = int(g["rows"]), int(g["cols"]), int(g["seed"])
    if rows == cols and rows >= 2048:
        pytest.skip("square 2048+ takes tens of seconds on one core; covered by make_golden --big cross-check")
    a = oracle_lib.generate(rows, cols, seed).astype(np.float64)
    out = oracle_lib.oracle_solve(np.asfortranarray(a))
    assert out["status"] == oracle_lib.ORC_OK
    # identical edges in the reference's own order, identical values
    assert np.array_equal(out["x"], g["x"]) and np.array_equal(out["y"], g["y"])
    assert float(out["v"].sum()) == float(g["objective"])
    assert out["stats"]["reverse_work"] == 0 and out["stats"]["lambda_rejects"] == 0


@pytest.mark.parametrize("path", [p for p in parity.golden_cases(max_rows=2048) if "xf" in np.load(p)])
def test_oracle_matches_golden_float(oracle_lib, path):
    g = np.load(path)
    a = oracle_lib.generate(int(g["rows"]), int(g["cols"]), int(g["seed"]))
    out = oracle_lib.oracle_solve(np.asfortranarray(a), max_bids=10_000_000)
    assert out["status"] == oracle_lib.ORC_OK
    assert np.array_equal(out["x"], g["xf"]) and np.array_equal(out["y"], g["yf"])


def test_oracle_layouts_agree(oracle_lib):
    a = oracle_lib.generate(200, 300, 7).astype(np.float64)
    r = oracle_lib.oracle_solve(np.ascontiguousarray(a))
    c = oracle_lib.oracle_solve(np.asfortranarray(a))
    assert np.array_equal(r["x"], c["x"]) and np.array_equal(r["y"], c["y"])


def test_oracle_matches_scipy_optimum(oracle_lib):
    from scipy.optimize import linear_sum_assignment
    for rows, cols, seed in [(32, 48, 11), (200, 300, 5), (128, 128, 4)]:
        a = oracle_lib.generate(rows, cols, seed).astype(np.float64)
        out = oracle_lib.oracle_solve(a)
        ri, ci = linear_sum_assignment(a, maximize=True)
        assert abs(out["v"].sum() - a[ri, ci].sum()) <= rows * 1e-5 / cols


def test_oracle_zero_means_no_edge(oracle_lib):
    # a == 0 is skipped (Auction.h:134): a banded matrix must be solved inside its band
    a = oracle_lib.generate(40, 60, 3).astype(np.float64)
    i, j = np.indices(a.shape)
    a[(j < i) | (j > i + 20)] = 0.0
    out = oracle_lib.oracle_solve(a)
    assert out["status"] == oracle_lib.ORC_OK and (out["v"] > 0).all()
    assert ((out["y"] >= out["x"]) & (out["y"] <= out["x"] + 20)).all()


def test_oracle_bid_budget_and_float_livelock(oracle_lib):
    # Auction<float> livelocks on square inputs of this size: budget must end it
    a = oracle_lib.generate(512, 512, 1)
    out = oracle_lib.oracle_solve(a, max_bids=300_000)
    assert out["status"] in (oracle_lib.ORC_BUDGET, oracle_lib.ORC_OK)
    a64 = oracle_lib.generate(64, 64, 1).astype(np.float64)
    out = oracle_lib.oracle_solve(a64, max_bids=10)
    assert out["status"] == oracle_lib.ORC_BUDGET and out["stats"]["bids"] == 10


def test_oracle_batch_matches_golden(oracle_lib):
    g = np.load(parity.GOLDEN_DIR + "/batch_256x32x48_s1.npz")
    batch, rows, cols, seed = (int(g[k]) for k in ("batch", "rows", "cols", "seed"))
    a = np.stack([oracle_lib.generate(rows, cols, seed, problem=b) for b in range(batch)])
    out, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=2)
    assert failed == 0 and np.array_equal(out, g["row_to_col"])
    outf, failed = oracle_lib.oracle_batch(a, use_double=False, n_threads=3)
    assert failed == 0 and np.array_equal(outf, g["row_to_col_f"])


def test_oracle_against_compiled_reference(oracle_lib):
    if oracle_lib.reference() is None:
        pytest.skip("oracle/_ref not built here")
    for rows, cols, seed in [(5, 9, 4), (48, 64, 9), (150, 150, 2), (300, 450, 6)]:
        a = np.asfortranarray(oracle_lib.generate(rows, cols, seed).astype(np.float64))
        o, r = oracle_lib.oracle_solve(a), oracle_lib.reference_solve(a)
        assert np.array_equal(o["x"], r["x"]) and np.array_equal(o["y"], r["y"]) and np.array_equal(o["v"], r["v"])


def test_reference_mt_is_wrong_on_asymmetric(oracle_lib):
    # AuctionMT only scans columns [0, rows) on dense rows < cols input
    if oracle_lib.reference() is None:
        pytest.skip("oracle/_ref not built here")
    a = np.asfortranarray(oracle_lib.generate(60, 90, 1).astype(np.float64))
    st, mt = oracle_lib.reference_solve(a), oracle_lib.reference_solve(a, "mt", 2)
    assert mt["y"].max() < 60 and st["v"].sum() > mt["v"].sum()
    sq = np.asfortranarray(oracle_lib.generate(60, 60, 1).astype(np.float64))
    st, mt = oracle_lib.reference_solve(sq), oracle_lib.reference_solve(sq, "mt", 2)
    assert np.array_equal(st["x"], mt["x"]) and np.array_equal(st["y"], mt["y"])
