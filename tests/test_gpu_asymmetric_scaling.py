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
@pytest.mark.parametrize("opt", [dict(eps_scaling=1), dict(eps_scaling=1, eps_start=0.25), dict(eps_scaling=1, row_cache=-1),
                                 dict(eps_scaling=2)], ids=["auto", "classic", "scanned", "last-phase-repair"])
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


def test_scaling_pays_where_the_fixed_eps_crawls(handle):
    # a = u_i * v_j (every row wants the same columns): prices must climb far above the spread of a row's best values and
    # a fixed eps = 1e-5 / cols crawls there.  With the repair at the end of every phase (eps_scaling = 1) the scaled
    # solve needs many times fewer rounds (measured at 2048 x 2052: 2 869 718 -> 114 197); with the repair after the last
    # phase only (eps_scaling = 2) nothing is gained — the last phase replays the whole price war at the final eps.
    rows, cols = 1024, 1028
    rng = np.random.default_rng(11)
    from test_gpu_structured import _rank_one
    a = np.ascontiguousarray(_rank_one(rng, rows, cols))
    eps = 1e-5 / cols
    fixed, pf = handle.solve(a, want_prices=True, max_rounds=50_000_000)
    fixed_rounds = handle.last_stats["rounds"]
    scaled, ps = handle.solve(a, want_prices=True, eps_scaling=1, max_rounds=50_000_000)
    st = dict(handle.last_stats)
    for r2c, p in ((fixed, pf), (scaled, ps)):
        assert parity.is_valid_assignment(r2c, cols) and _post_condition(r2c, p)
        assert parity.eps_cs_violation(a, r2c, p) <= eps + 8 * np.spacing(max(1.0, float(p.max())))
    assert abs(parity.objective(a, scaled) - parity.objective(a, fixed)) <= rows * eps + 1e-9
    assert st["warm_fallback"] >= 1
    assert st["rounds"] * 5 <= fixed_rounds, (st["rounds"], fixed_rounds)


def test_scaled_uniform_near_square_problem_costs_a_bounded_factor(handle, oracle_lib):
    # where the fixed eps does not crawl, the per-phase repair costs rounds (measured ~3x on 2048 x 2052 uniform noise)
    rows, cols = 1024, 1028
    a = oracle_lib.generate(rows, cols, 3)
    eps = 1e-5 / cols
    fixed = handle.solve(a, max_rounds=50_000_000)
    fixed_rounds = handle.last_stats["rounds"]
    for mode, factor in ((1, 6.0), (2, 1.5)):
        scaled, ps = handle.solve(a, want_prices=True, eps_scaling=mode, max_rounds=50_000_000)
        assert np.array_equal(scaled, fixed) and _post_condition(scaled, ps)
        assert handle.last_stats["rounds"] <= factor * fixed_rounds, (mode, handle.last_stats["rounds"], fixed_rounds)
