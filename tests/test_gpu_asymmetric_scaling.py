"""eps scaling on rows < cols: the forward auction plus the repair that does the reverse
auction's job (Auction.h:346-469: lambda raised until no unassigned object is priced above an assigned one;
AuctionCommon.h:347-354) — here a floor raise that keeps prices monotone.  Checked against the reference (goldens made
by the unmodified reference, which runs at the fixed eps) and against the CPU port on structured near-square inputs."""
import numpy as np
import pytest

import parity
from test_gpu_structured import GENERATORS

pytestmark = pytest.mark.gpu


def _post_condition(r2c, prices):
    free = np.ones(len(prices), bool)
    free[r2c] = False
    return (not free.any()) or prices[free].max() <= prices[~free].min()


@pytest.mark.parametrize("name", ["auction_3x7_s2", "auction_8x12_s1", "auction_32x48_s3", "auction_100x101_s1", "auction_200x300_s1",
                                  "auction_200x300_s3", "auction_511x513_s1", "auction_512x768_s1", "auction_1024x1536_s1",
                                  "auction_2048x3072_s2", "auction_8192x12288_s1"])
@pytest.mark.parametrize("opt", [dict(eps_scaling=1), dict(eps_scaling=1, eps_start=0.25), dict(eps_scaling=1, row_cache=-1)],
                         ids=["auto", "classic", "scanned"])
def test_scaled_asymmetric_solve_matches_reference_golden(handle, oracle_lib, name, opt):
    g = np.load(f"{parity.GOLDEN_DIR}/{name}.npz")
    rows, cols, seed = int(g["rows"]), int(g["cols"]), int(g["seed"])
    if rows >= 8192 and opt.get("row_cache") == -1:
        pytest.skip("covered by the cached path")
    a = oracle_lib.generate(rows, cols, seed)
    r2c, prices = handle.solve(a, want_prices=True, **opt)
    st = handle.last_stats
    assert st["phases"] > 1
    assert parity.check_parity(a, r2c, prices, parity.golden_row_to_col(g), 1e-5 / cols) == "exact"
    assert _post_condition(r2c, prices)          # AuctionCommon.h:347-354 with lambda = lowest assigned price


@pytest.mark.parametrize("gen", GENERATORS, ids=lambda g: g.__name__.strip("_"))
@pytest.mark.parametrize("shape", [(150, 230), (250, 256), (1000, 1004)])
def test_scaled_asymmetric_structured_inputs(handle, oracle_lib, gen, shape):
    rows, cols = shape
    rng = np.random.default_rng(77 + rows)
    a = np.ascontiguousarray(gen(rng, rows, cols))
    eps = 1e-5 / cols
    r2c, prices = handle.solve(a, want_prices=True, eps_scaling=1, max_rounds=20_000_000)
    st = dict(handle.last_stats)
    assert parity.is_valid_assignment(r2c, cols)
    tol = eps + 8 * np.spacing(max(1.0, float(prices.max())))
    assert parity.eps_cs_violation(a, r2c, prices) <= tol
    assert _post_condition(r2c, prices)
    # the reference's own schedule (fixed eps, forward only) on the same input: within rows * eps of each other
    fixed, pf = handle.solve(a, want_prices=True, max_rounds=20_000_000)
    assert abs(parity.objective(a, r2c) - parity.objective(a, fixed)) <= rows * eps + 1e-9
    if rows <= 250:
        out = oracle_lib.oracle_solve(a.astype(np.float64), max_bids=30_000_000)
        if out["status"] == 0 and len(out["x"]) == rows:
            ref = np.full(rows, -1, np.int64)
            ref[out["x"]] = out["y"]
            parity.check_parity(a, r2c, prices, ref, eps)


def test_scaled_near_square_problem_is_as_good_and_not_slower_by_much(handle):
    # 4096 x 4100, two clusters.  Measured (tools/asym_pay.py, B200): fixed eps 70 814 rounds, scaled 73 184 with one
    # repair — the floor raise after the LAST phase re-plays at the final eps the price war that the fixed-eps solve
    # plays, so scaling does not pay yet on near-square asymmetric inputs (it would with the repair run at the end of
    # every phase, at that phase's eps: DESIGN.md section 9).  What must hold: the same quality of answer.
    rows, cols = 4096, 4100
    rng = np.random.default_rng(11)
    from test_gpu_structured import _two_clusters
    a = np.ascontiguousarray(_two_clusters(rng, rows, cols))
    eps = 1e-5 / cols
    fixed, pf = handle.solve(a, want_prices=True, max_rounds=50_000_000)
    fixed_rounds = handle.last_stats["rounds"]
    scaled, ps = handle.solve(a, want_prices=True, eps_scaling=1, max_rounds=50_000_000)
    st = handle.last_stats
    assert parity.is_valid_assignment(scaled, cols) and _post_condition(scaled, ps)
    assert parity.eps_cs_violation(a, scaled, ps) <= eps + 8 * np.spacing(max(1.0, float(ps.max())))
    assert abs(parity.objective(a, scaled) - parity.objective(a, fixed)) <= rows * eps + 1e-9
    assert st["rounds"] <= 1.5 * fixed_rounds, (st["rounds"], fixed_rounds)
