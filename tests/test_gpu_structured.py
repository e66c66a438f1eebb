"""GPU parity on inputs that are NOT uniform noise: the first eps of the scaling schedule is sized from
the data (k_probe), and the row caches rest on a float filter, so both must hold on structured
matrices too.  Each case is checked against the CPU oracle with the parity rule of SURVEY.md 8(c);
the oracle runs with a bid budget because the reference's fixed eps is slow on some structures."""
import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu


def _geometric(rng, rows, cols):
    # MHT-like gating: benefit = 1 - distance between track and measurement positions in the unit square
    t = rng.random((rows, 2))
    m = rng.random((cols, 2))
    d = np.sqrt(((t[:, None, :] - m[None, :, :]) ** 2).sum(-1)) / np.sqrt(2.0)
    return (1.0 - d).astype(np.float32)


def _rank_one(rng, rows, cols):
    u = 0.2 + 0.8 * rng.random(rows)
    v = 0.2 + 0.8 * rng.random(cols)
    return (np.outer(u, v) * (1.0 + 1e-3 * rng.random((rows, cols)))).astype(np.float32).clip(1e-6, 1.0)


def _heavy_tail(rng, rows, cols):
    return (rng.random((rows, cols)) ** 8).astype(np.float32).clip(1e-7, 1.0)


def _narrow_band(rng, rows, cols):
    return (0.5 + 1e-3 * rng.random((rows, cols))).astype(np.float32)


def _banded_sparse(rng, rows, cols):
    a = rng.random((rows, cols)).astype(np.float32).clip(1e-6, 1.0)
    i, j = np.indices(a.shape)
    a[(j < i - 3) | (j > i + cols // 4)] = 0.0
    return a


def _two_clusters(rng, rows, cols):
    a = rng.random((rows, cols)).astype(np.float32) * 0.1
    a[: rows // 2, : cols // 2] += 0.9
    a[rows // 2:, cols // 2:] += 0.9
    return a.clip(1e-6, 1.0)


GENERATORS = [_geometric, _rank_one, _heavy_tail, _narrow_band, _banded_sparse, _two_clusters]


@pytest.mark.parametrize("gen", GENERATORS, ids=lambda g: g.__name__.strip("_"))
@pytest.mark.parametrize("shape", [(96, 96), (200, 200), (150, 230)])
def test_structured_inputs_match_the_oracle(handle, oracle_lib, gen, shape):
    rng = np.random.default_rng(1234 + shape[0])
    a = np.ascontiguousarray(gen(rng, *shape))
    out = oracle_lib.oracle_solve(a.astype(np.float64), max_bids=30_000_000)
    if out["status"] != 0 or len(out["x"]) != shape[0]:
        pytest.skip("the reference itself does not finish on this input within the bid budget")
    ref = np.full(shape[0], -1, np.int64)
    ref[out["x"]] = out["y"]
    eps = 1e-5 / shape[1]
    r2c, prices = handle.solve(a, want_prices=True)
    parity.check_parity(a, r2c, prices, ref, eps)
    # the row-scanning path and the classic schedule must agree with it on the objective as well
    r2c2, prices2 = handle.solve(a, want_prices=True, row_cache=-1, eps_start=0.25)
    parity.check_parity(a, r2c2, prices2, ref, eps)


@pytest.mark.parametrize("gen", GENERATORS, ids=lambda g: g.__name__.strip("_"))
def test_structured_1024_cached_and_scanned_paths_agree_bit_for_bit(handle, gen):
    """Same schedule, row caches on and off: identical prices (the cache only ever answers exactly).
    The data-sized schedule (which may restart from a larger eps on such inputs) must end eps-CS too."""
    rng = np.random.default_rng(99)
    a = np.ascontiguousarray(gen(rng, 1024, 1024))
    eps = 1e-5 / 1024

    def cs_ok(r2c, prices):
        # a - p is rounded at the magnitude of the prices (rows with few edges drive them far above 1)
        tol = eps + 8 * np.spacing(max(1.0, float(prices.max())))
        return parity.is_valid_assignment(r2c, 1024) and parity.eps_cs_violation(a, r2c, prices) <= tol

    r2c0, prices0 = handle.solve(a, want_prices=True)                      # data-sized schedule
    assert cs_ok(r2c0, prices0)
    r2c, prices = handle.solve(a, want_prices=True, eps_start=0.02)       # row caches
    assert handle.last_stats["tail_runs"] == 0                            # an explicit start is never restarted
    r2c2, prices2 = handle.solve(a, want_prices=True, eps_start=0.02, row_cache=-1)
    assert handle.last_stats["cache_hits"] == 0
    assert np.array_equal(r2c, r2c2) and np.array_equal(prices, prices2)
    assert cs_ok(r2c, prices)
    # all three end within rows * eps of each other (each is within rows * eps of the optimum)
    assert abs(parity.objective(a, r2c0) - parity.objective(a, r2c)) <= 1024 * eps + 1e-9


def test_too_small_a_start_is_restarted_from_a_larger_eps(handle):
    """a = u_i * v_j: every row wants the same columns, equilibrium prices are far above the spread of a
    row's best values that the probe measures; the first phase runs over its budget and the schedule
    restarts from eps * factor^2 (escalate_schedule)."""
    rng = np.random.default_rng(5)
    a = np.ascontiguousarray(_rank_one(rng, 1024, 1024))
    r2c, prices = handle.solve(a, want_prices=True)
    st = handle.last_stats
    assert st["tail_runs"] >= 1 and st["cache_hits"] > 0
    assert parity.is_valid_assignment(r2c, 1024)
    assert parity.eps_cs_violation(a, r2c, prices) <= 1e-5 / 1024 + 8 * np.spacing(max(1.0, float(prices.max())))
