"""GPU tests of the batched mode (MHT-style 32x48 problems) and of the device-resident entry
points, through the C-ABI."""
import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu


def _batch_inputs(oracle_lib, batch, rows, cols, seed, problem0=0):
    return np.stack([oracle_lib.generate(rows, cols, seed, problem=problem0 + b) for b in range(batch)])


@pytest.mark.parametrize("name", ["batch_256x32x48_s1.npz", "batch_64x8x12_s1.npz"])
def test_batch_matches_reference_exactly(handle, oracle_lib, name):
    # the batched kernel keeps the reference's Gauss-Seidel order: identical assignment, always
    g = np.load(parity.GOLDEN_DIR + "/" + name)
    batch, rows, cols, seed = (int(g[k]) for k in ("batch", "rows", "cols", "seed"))
    a = _batch_inputs(oracle_lib, batch, rows, cols, seed)
    out = handle.solve_batch(a)
    assert np.array_equal(out, g["row_to_col"])
    a_cm = np.ascontiguousarray(a.transpose(0, 2, 1))  # each problem column-major, Eigen's default
    import auctionalgorithm_b200 as ab
    out_cm = handle.solve_batch(a_cm.reshape(batch, rows, cols), layout=ab.COL_MAJOR)
    assert np.array_equal(out_cm, g["row_to_col"])


@pytest.mark.parametrize("shape", [(1, 1), (1, 7), (5, 5), (17, 33), (32, 32), (32, 64), (31, 63), (20, 21)])
def test_batch_shapes_against_oracle(handle, oracle_lib, shape):
    rows, cols = shape
    a = _batch_inputs(oracle_lib, 200, rows, cols, 9)
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    assert np.array_equal(handle.solve_batch(a), ref)


def test_batch_with_zeros_and_ties(handle, oracle_lib):
    a = _batch_inputs(oracle_lib, 300, 16, 24, 4)
    rng = np.random.default_rng(0)
    a[rng.random(a.shape) < 0.3] = 0.0
    a[:, :, :2] = np.maximum(a[:, :, :2], 0.25)       # keep every row feasible
    a[:50] = np.round(a[:50] * 8) / 8                 # many exact ties
    a[:50, :, :2] = np.maximum(a[:50, :, :2], 0.125)
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    assert np.array_equal(handle.solve_batch(a), ref)


# ---- the filtered kernel (cols % 4 == 0): every way its single-precision pass can be wrong-footed -------------
@pytest.mark.parametrize("shape", [(1, 4), (3, 4), (4, 4), (8, 8), (7, 12), (20, 20), (32, 36), (32, 44), (30, 60),
                                   (32, 48), (16, 56), (2, 64)])
def test_filtered_batch_shapes_against_oracle(handle, oracle_lib, shape):
    # whole and half last chunks (cols % 8 == 0 / 4), batches that do not fill a warp's sixteen slots
    rows, cols = shape
    for batch in (1, 15, 333):
        a = _batch_inputs(oracle_lib, batch, rows, cols, 21)
        ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
        assert failed == 0
        assert np.array_equal(handle.solve_batch(a), ref)


def test_filtered_batch_near_ties(handle, oracle_lib):
    # leaders one to a few fp32 ulps apart, and exact ties: the fp32 pass cannot order them, the bound has to
    # send those bids to the exact scan (and the lowest column has to win the exact ties)
    rng = np.random.default_rng(5)
    a = _batch_inputs(oracle_lib, 400, 24, 32, 8)
    base = np.float32(0.75)
    for b in range(400):
        cols = rng.choice(32, size=6, replace=False)
        ulps = rng.integers(0, 4, size=(24, 6))
        a[b][:, cols] = (base + ulps * np.spacing(base)).astype(np.float32)
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    assert np.array_equal(handle.solve_batch(a), ref)


def test_filtered_batch_coarse_grid_many_ties(handle, oracle_lib):
    a = _batch_inputs(oracle_lib, 500, 32, 48, 13)
    a = (np.ceil(a * 4) / 4).astype(np.float32)          # four distinct values: ties everywhere
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    assert np.array_equal(handle.solve_batch(a), ref)


def test_filtered_batch_sparse_rows_and_single_edges(handle, oracle_lib):
    # most coefficients are "no edge"; some rows keep a single edge (second best = -1e6, Auction.h:237):
    # zero leaders switch the warp to the zero-aware pass, single edges always take the exact scan
    rng = np.random.default_rng(6)
    a = _batch_inputs(oracle_lib, 600, 12, 24, 17)
    a[rng.random(a.shape) < 0.8] = 0.0
    for r in range(12):
        a[:, r, r] = np.maximum(a[:, r, r], 0.05)          # feasible: the diagonal is always an edge
    a[::3, 0, 1:] = 0.0                                     # row 0 of every third problem: one edge only
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    assert np.array_equal(handle.solve_batch(a), ref)


def test_filtered_batch_wide_magnitudes(handle, oracle_lib):
    # coefficients outside [0, 1]: large values make large prices (the bound scales with the price), tiny values
    # sit next to them, negative values are legal edges
    rng = np.random.default_rng(7)
    a = _batch_inputs(oracle_lib, 300, 16, 24, 19)
    a[:100] *= np.float32(1000.0)
    a[100:200] = (a[100:200] - np.float32(0.5)) * np.float32(4.0)
    a[200:] *= (10.0 ** rng.integers(-6, 3, size=a[200:].shape)).astype(np.float32)
    a[a == 0] = np.float32(0.5)
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    assert np.array_equal(handle.solve_batch(a), ref)


def test_filtered_batch_large_batch_is_consistent(handle, oracle_lib):
    # more problems than one resident wave holds slots for: every warp hands its slice out in turn; sampled
    # against the oracle, and the whole batch against a second run split in two
    import torch
    n = 50000
    t = torch.empty((n, 32, 48), dtype=torch.float32, device="cuda")
    handle.generate_device(t.data_ptr(), n, 32, 48, 48, 1, seed=3, problem0=0)
    out = torch.empty((n, 32), dtype=torch.int32, device="cuda")
    handle.solve_batch_device(t.data_ptr(), n, 32, 48, out.data_ptr())
    got = out.cpu().numpy()
    assert (np.sort(got, axis=1)[:, 1:] != np.sort(got, axis=1)[:, :-1]).all() and got.min() >= 0 and got.max() < 48
    idx = np.concatenate([np.arange(0, 64), np.arange(n - 64, n), np.arange(777, n, 997)])
    a = np.stack([oracle_lib.generate(32, 48, 3, problem=int(b)) for b in idx])
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0 and np.array_equal(got[idx], ref)
    out2 = torch.empty((n, 32), dtype=torch.int32, device="cuda")
    h = 20011
    handle.solve_batch_device(t.data_ptr(), h, 32, 48, out2.data_ptr())
    handle.solve_batch_device(t[h:].data_ptr(), n - h, 32, 48, out2[h:].data_ptr())
    assert torch.equal(out, out2)


# ---- bid cap and whole-warp takeover (batch_warp_finish), dynamic hand-out ------------------------------------
def _oracle_bids(oracle_lib, a):
    return [oracle_lib.oracle_solve(p.astype(np.float64))["stats"]["bids"] for p in a]


def test_filtered_batch_long_price_wars_of_the_bench_shards(handle, oracle_lib):
    # the two slowest problems of the bench's seed-1 shards (1483 and 2148 bids; the mean is 44) among neighbours:
    # they reach the default cap of 96 bids and are finished by a whole warp, continuing the same sweep
    ids = [1117884, 3097078, 274277, 3248947] + list(range(1117884 - 20, 1117884 + 20))
    a = np.stack([oracle_lib.generate(32, 48, 1, problem=i) for i in ids])
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    got = handle.solve_batch(a)
    assert np.array_equal(got, ref)
    bids = _oracle_bids(oracle_lib, a[:4])
    assert bids[0] == 1483 and bids[1] == 2148
    st = handle.last_stats
    assert st["batch_max_bids"] == 2148 and st["batch_takeovers"] >= 4
    # without the cap: same answer from the two-lane pass alone
    assert np.array_equal(handle.solve_batch(a, batch_bid_cap=-1), ref)
    assert handle.last_stats["batch_takeovers"] == 0 and handle.last_stats["batch_max_bids"] == 2148


@pytest.mark.parametrize("shape", [(32, 48), (8, 12), (32, 64), (20, 44), (31, 36), (5, 8), (1, 4)])
@pytest.mark.parametrize("cap", [1, 8])
def test_filtered_batch_takeover_forced(handle, oracle_lib, shape, cap):
    # a cap of 8 (or 1) bids sends every problem through the takeover after its first bids: the warp must continue
    # exactly where the two lanes stopped, for every shape (half last chunks, fewer than 32 / more than 32 columns)
    rows, cols = shape
    for batch in (1, 40, 700):
        a = _batch_inputs(oracle_lib, batch, rows, cols, 31)
        ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
        assert failed == 0
        assert np.array_equal(handle.solve_batch(a, batch_bid_cap=cap), ref)
        if rows > cap:
            assert handle.last_stats["batch_takeovers"] == batch


def test_filtered_batch_takeover_with_zeros_ties_and_single_edges(handle, oracle_lib):
    rng = np.random.default_rng(11)
    a = _batch_inputs(oracle_lib, 500, 16, 24, 23)
    a[rng.random(a.shape) < 0.7] = 0.0
    for r in range(16):
        a[:, r, r] = np.maximum(a[:, r, r], 0.05)
    a[::3, 0, 1:] = 0.0                                  # single-edge rows: second best = -1e6
    a[100:200] = (np.ceil(a[100:200] * 4) / 4).astype(np.float32)   # coarse grid: ties everywhere
    for r in range(16):
        a[100:200, r, r] = np.maximum(a[100:200, r, r], 0.25)
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    for cap in (2, 5, 0):
        assert np.array_equal(handle.solve_batch(a, batch_bid_cap=cap), ref)


def test_filtered_batch_synthetic_price_war(handle, oracle_lib):
    # k + 1 rows whose k best columns coincide with equal values, every other column delta lower: the rows outbid
    # each other in steps of eps until the prices have risen by delta, about k * delta / eps bids (30 000 here)
    rows, cols, k = 32, 48, 4
    eps = 1e-5 / cols
    a = _batch_inputs(oracle_lib, 24, rows, cols, 41) * np.float32(0.5)
    delta = np.float32(50000 * eps / (k + 1))
    for b in range(24):
        a[b, :k + 1, :] = np.float32(0.9) - delta
        a[b, :k + 1, 7:7 + k] = np.float32(0.9)
        a[b, k + 1:, 7:7 + k] = np.float32(0.001)       # the other rows stay out of the war
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0
    bids = _oracle_bids(oracle_lib, a[:1])[0]
    assert 20000 < bids < 200000
    got = handle.solve_batch(a)
    assert np.array_equal(got, ref)
    assert handle.last_stats["batch_max_bids"] >= bids and handle.last_stats["batch_takeovers"] == 24
    # the watchdog still counts across the takeover
    import auctionalgorithm_b200 as ab
    with pytest.raises(ab.B200LsapError) as e:
        handle.solve_batch(a, max_rounds=5000)
    assert e.value.status == ab.ERR_MAX_ROUNDS


@pytest.mark.parametrize("batch", [1, 2, 16, 17, 63, 64, 65, 1000, 4099, 37889, 80000])
def test_filtered_batch_dynamic_handout_covers_every_problem(handle, oracle_lib, batch):
    # reservations of 16 problems from one counter: every problem solved exactly once whatever the batch size
    # (partial last reservation, fewer problems than warps, more than one resident wave)
    import torch
    t = torch.empty((batch, 32, 48), dtype=torch.float32, device="cuda")
    handle.generate_device(t.data_ptr(), batch, 32, 48, 48, 1, seed=7, problem0=5)
    out = torch.full((batch, 32), -7, dtype=torch.int32, device="cuda")
    handle.solve_batch_device(t.data_ptr(), batch, 32, 48, out.data_ptr())
    got = out.cpu().numpy()
    s = np.sort(got, axis=1)
    assert got.min() >= 0 and got.max() < 48 and (s[:, 1:] != s[:, :-1]).all()
    idx = np.unique(np.concatenate([np.arange(0, min(batch, 40)), np.arange(max(0, batch - 40), batch),
                                    np.arange(0, batch, 211)]))
    a = np.stack([oracle_lib.generate(32, 48, 7, problem=5 + int(b)) for b in idx])
    ref, failed = oracle_lib.oracle_batch(a, use_double=True, n_threads=4)
    assert failed == 0 and np.array_equal(got[idx], ref)


def test_batch_errors(handle):
    import auctionalgorithm_b200 as ab
    with pytest.raises(ab.B200LsapError) as e:
        handle.solve_batch(np.ones((2, 33, 40), np.float32))
    assert e.value.status == ab.ERR_SHAPE
    a = np.ones((3, 4, 6), np.float32)
    a[1, 2] = 0
    with pytest.raises(ab.B200LsapError) as e:
        handle.solve_batch(a)
    assert e.value.status == ab.ERR_INFEASIBLE
    assert handle.solve_batch(np.ones((0, 4, 6), np.float32)).shape == (0, 4)


def test_device_generator_is_bit_identical(handle, oracle_lib):
    import torch
    import auctionalgorithm_b200 as ab
    for batch, rows, cols, ld, layout in [(3, 32, 48, 48, ab.ROW_MAJOR), (1, 100, 130, 160, ab.ROW_MAJOR),
                                           (2, 20, 30, 20, ab.COL_MAJOR)]:
        major = rows if layout == ab.ROW_MAJOR else cols
        minor = cols if layout == ab.ROW_MAJOR else rows
        t = torch.full((batch, major, ld), -1.0, dtype=torch.float32, device="cuda")
        handle.generate_device(t.data_ptr(), batch, rows, cols, ld, layout, seed=5, problem0=7)
        got = t.cpu().numpy()
        for b in range(batch):
            want = oracle_lib.generate(rows, cols, 5, problem=7 + b)
            blk = got[b, :, :minor]
            assert np.array_equal(blk if layout == ab.ROW_MAJOR else blk.T, want)
            assert (got[b, :, minor:] == 0).all()
    # column shard of a bigger problem
    t = torch.empty((64, 32), dtype=torch.float32, device="cuda")
    handle.generate_shard_device(t.data_ptr(), 64, 32, 32, col0=40, cols_global=100, seed=3)
    assert np.array_equal(t.cpu().numpy(), oracle_lib.generate(64, 100, 3)[:, 40:72])


def test_device_solve_matches_host_solve(handle, oracle_lib):
    import torch
    import auctionalgorithm_b200 as ab
    for rows, cols in [(200, 300), (256, 256), (123, 457)]:
        a = oracle_lib.generate(rows, cols, 12)
        want, wprices = handle.solve(a, want_prices=True)
        for layout, host in ((ab.ROW_MAJOR, a), (ab.COL_MAJOR, np.asfortranarray(a))):
            t = torch.from_numpy(np.ascontiguousarray(host if layout == ab.ROW_MAJOR else host.T)).cuda()
            r2c = torch.empty(rows, dtype=torch.int32, device="cuda")
            prices = torch.empty(cols, dtype=torch.float64, device="cuda")
            handle.solve_device(t.data_ptr(), rows, cols, cols if layout == ab.ROW_MAJOR else rows, layout,
                                r2c.data_ptr(), prices.data_ptr())
            assert np.array_equal(r2c.cpu().numpy(), want) and np.array_equal(prices.cpu().numpy(), wprices)


def test_batch_device_and_stream_handoff(handle, oracle_lib):
    import torch
    g = np.load(parity.GOLDEN_DIR + "/batch_256x32x48_s1.npz")
    t = torch.empty((256, 32, 48), dtype=torch.float32, device="cuda")
    handle.generate_device(t.data_ptr(), 256, 32, 48, 48, 1, seed=1, problem0=0)
    out = torch.empty((256, 32), dtype=torch.int32, device="cuda")
    handle.solve_batch_device(t.data_ptr(), 256, 32, 48, out.data_ptr())
    assert np.array_equal(out.cpu().numpy(), g["row_to_col"])


def test_bid_sweep_reports_bytes(handle, oracle_lib):
    import torch
    t = torch.empty((1024, 2048), dtype=torch.float32, device="cuda")
    handle.generate_device(t.data_ptr(), 1, 1024, 2048, 2048, 1, seed=2)
    ms, nbytes = handle.bid_sweep_device(t.data_ptr(), 1024, 2048, 2048, use_last_prices=False)
    assert nbytes == 1024 * 2048 * 4 and ms > 0
