"""GPU tests of the column-sharded solve (b200lsap_shard_*): several shards emulated on one GPU
(one handle per shard, the all-gather replaced by a concatenation), against the reference-made
golden vectors and against the single-GPU solve."""
import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu


def _run(oracle_lib, rows, cols, seed, world, want_prices=True, a=None, **options):
    import torch
    import auctionalgorithm_b200 as ab
    from auctionalgorithm_b200 import sharded
    if a is None:
        a = oracle_lib.generate(rows, cols, seed)
    handles = [ab.Handle(0) for _ in range(world)]
    engines = []
    for h, (c0, cl) in zip(handles, sharded.split_columns(cols, world)):
        ld = (cl + 3) // 4 * 4                                    # zero columns up to a multiple of 4
        local = np.zeros((rows, ld), np.float32)
        local[:, :cl] = a[:, c0:c0 + cl]
        t = torch.from_numpy(local).cuda()
        engines.append(sharded.CudaShardEngine(h, t, rows, cl, ld, c0, cols, **options))
    res = sharded.solve_sharded_emulated(engines, rows, want_prices=want_prices)
    for h in handles:
        h.close()
    return a, res


@pytest.mark.parametrize("name,world", [("auction_200x300_s1", 2), ("auction_200x300_s2", 3), ("auction_2048x3072_s1", 4),
                                        ("auction_1024x1536_s1", 8), ("auction_256x256_s2", 2), ("auction_511x513_s1", 3),
                                        ("auction_100x101_s1", 2)])
def test_sharded_matches_golden_and_replicas_agree(handle, oracle_lib, name, world):
    g = np.load(f"{parity.GOLDEN_DIR}/{name}.npz")
    rows, cols, seed = int(g["rows"]), int(g["cols"]), int(g["seed"])
    a, res = _run(oracle_lib, rows, cols, seed, world)
    r2c0, p0, st0 = res[0]
    for r2c, p, st in res[1:]:
        assert bool((r2c == r2c0).all()) and bool((p == p0).all()) and st["rounds"] == st0["rounds"]
    r2c = r2c0.cpu().numpy().astype(np.int64)
    verdict = parity.check_parity(a, r2c, p0.cpu().numpy(), parity.golden_row_to_col(g), 1e-5 / cols)
    assert verdict == "exact"
    # the sharded rounds are the wide (Jacobi) rounds of the single-GPU solver: same prices, bit for bit
    ref_r2c, ref_p = handle.solve(a, want_prices=True, tail_rows=-1, narrow_rows=-1, eps_start=0.25)  # the shards' schedule
    assert np.array_equal(ref_r2c, r2c) and np.array_equal(ref_p, p0.cpu().numpy())
    assert handle.last_stats["rounds"] == st0["rounds"] and handle.last_stats["bids"] == st0["bids"]


def test_sharded_host_driven_on_a_non_default_torch_stream(oracle_lib):
    # the exchange (here torch.cat) and the handle's kernels must share one stream: the engine binds the handle to
    # torch's CURRENT stream when it is made (the tests used to depend on launch timing)
    import torch
    g = np.load(f"{parity.GOLDEN_DIR}/auction_512x768_s1.npz")
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        a, res = _run(oracle_lib, 512, 768, 1, 3)
    side.synchronize()
    for r2c, p, st in res:
        assert np.array_equal(r2c.cpu().numpy().astype(np.int64), parity.golden_row_to_col(g))


def test_sharded_with_zeros_and_a_shard_without_edges(oracle_lib):
    rows, cols = 64, 512
    a = oracle_lib.generate(rows, cols, 11)
    a[:, 128:256] = 0.0          # rank 1 of 4 holds no edge at all
    a[::3, :100] = 0.0
    out = oracle_lib.oracle_solve(a.astype(np.float64))
    ref = np.full(rows, -1, np.int64)
    ref[out["x"]] = out["y"]
    _, res = _run(oracle_lib, rows, cols, 0, 4, a=a)
    r2c, p, _ = res[0]
    parity.check_parity(a, r2c.cpu().numpy().astype(np.int64), p.cpu().numpy(), ref, 1e-5 / cols)


def test_sharded_reports_infeasible_rows(oracle_lib):
    import auctionalgorithm_b200 as ab
    a = oracle_lib.generate(8, 256, 2)
    a[3, :] = 0.0
    with pytest.raises(ab.B200LsapError) as e:
        _run(oracle_lib, 8, 256, 0, 2, a=a)
    assert e.value.status == ab.ERR_INFEASIBLE


def test_shard_argument_checks(handle):
    import torch
    import auctionalgorithm_b200 as ab
    t = torch.zeros((8, 64), dtype=torch.float32, device="cuda")
    with pytest.raises(ab.B200LsapError):
        handle.shard_begin(t.data_ptr(), 8, 64, 64, 2, 128)      # col0 not a multiple of 4
    with pytest.raises(ab.B200LsapError):
        handle.shard_begin(t.data_ptr(), 8, 64, 64, 96, 128)     # shard runs past the last column
    with pytest.raises(ab.B200LsapError):
        handle.shard_bid(t.data_ptr())                           # no solve in progress


# ---- fused form: the exchange inside one persistent kernel per rank (b200lsap_solve_sharded_*) --------------------
def _run_fused(oracle_lib, rows, cols, seed, world, a=None, rows_cap=None, handles=None, **options):
    """`world` ranks on ONE GPU: a handle (own stream) per rank, mailboxes wired by pointer, every rank's kernel on
    148 / world SMs so that they are resident together."""
    import torch
    import auctionalgorithm_b200 as ab
    from auctionalgorithm_b200 import sharded
    if a is None:
        a = oracle_lib.generate(rows, cols, seed)
    own = handles is None
    if own:
        handles = [ab.Handle(0) for _ in range(world)]
        sharded.connect_fused_local(handles, rows_cap or rows)
    parts = sharded.split_columns(cols, world)
    shards, lds = [], []
    for c0, cl in parts:
        ld = (cl + 3) // 4 * 4
        local = np.zeros((rows, ld), np.float32)
        local[:, :cl] = a[:, c0:c0 + cl]
        shards.append(torch.from_numpy(local).cuda())
        lds.append(ld)
    torch.cuda.synchronize()
    try:
        res = sharded.solve_fused_local(handles, shards, rows, parts, lds, cols, want_prices=True, **options)
    finally:
        if own:
            for h in handles:
                h.close()
    return a, res


@pytest.mark.parametrize("name,world", [("auction_200x300_s1", 2), ("auction_200x300_s2", 3), ("auction_2048x3072_s1", 4),
                                        ("auction_1024x1536_s1", 8), ("auction_256x256_s2", 2), ("auction_511x513_s1", 3),
                                        ("auction_100x101_s1", 2), ("auction_32x48_s1", 1), ("auction_2048x2048_s1", 2)])
def test_fused_sharded_matches_golden_dense_solver_and_host_driven_form(handle, oracle_lib, name, world):
    g = np.load(f"{parity.GOLDEN_DIR}/{name}.npz")
    rows, cols, seed = int(g["rows"]), int(g["cols"]), int(g["seed"])
    a, res = _run_fused(oracle_lib, rows, cols, seed, world)
    r2c0, p0, st0 = res[0]
    for r2c, p, st in res[1:]:   # replicas: identical assignment, prices bit for bit, same rounds
        assert bool((r2c == r2c0).all()) and bool((p == p0).all()) and st["rounds"] == st0["rounds"]
    r2c = r2c0.cpu().numpy().astype(np.int64)
    assert parity.check_parity(a, r2c, p0.cpu().numpy(), parity.golden_row_to_col(g), 1e-5 / cols) == "exact"
    # the fused rounds are the wide (Jacobi) rounds of the single-GPU solver: same prices, bit for bit
    ref_r2c, ref_p = handle.solve(a, want_prices=True, tail_rows=-1, narrow_rows=-1, eps_start=0.25)
    assert np.array_equal(ref_r2c, r2c) and np.array_equal(ref_p, p0.cpu().numpy())
    assert handle.last_stats["rounds"] == st0["rounds"] and handle.last_stats["bids"] == st0["bids"]
    if world > 1 and rows <= 1024:
        _, host = _run(oracle_lib, rows, cols, seed, world)
        assert bool((host[0][0] == r2c0).all()) and bool((host[0][1] == p0).all())


def test_fused_sharded_mailboxes_are_reused_across_solves(oracle_lib):
    # flags count exchanges monotonically across solves (nothing is reset in between): different problems, fewer
    # rows than the mailboxes were created for, several solves on the same wiring
    import auctionalgorithm_b200 as ab
    from auctionalgorithm_b200 import sharded
    world = 3
    handles = [ab.Handle(0) for _ in range(world)]
    sharded.connect_fused_local(handles, 600)
    try:
        for rows, cols, seed in [(200, 300, 1), (512, 768, 1), (200, 300, 2), (64, 64, 3), (512, 768, 1)]:
            g = np.load(f"{parity.GOLDEN_DIR}/auction_{rows}x{cols}_s{seed}.npz")
            a, res = _run_fused(oracle_lib, rows, cols, seed, world, handles=handles)
            for r2c, p, st in res:
                assert np.array_equal(r2c.cpu().numpy().astype(np.int64), parity.golden_row_to_col(g))
    finally:
        for h in handles:
            h.close()


def test_fused_sharded_zero_columns_infeasible_rows_and_argument_checks(oracle_lib):
    import torch
    import auctionalgorithm_b200 as ab
    rows, cols = 64, 512
    a = oracle_lib.generate(rows, cols, 11)
    a[:, 128:256] = 0.0          # rank 1 of 4 holds no edge at all
    a[::3, :100] = 0.0
    out = oracle_lib.oracle_solve(a.astype(np.float64))
    ref = np.full(rows, -1, np.int64)
    ref[out["x"]] = out["y"]
    _, res = _run_fused(oracle_lib, rows, cols, 0, 4, a=a)
    r2c, p, _ = res[0]
    assert parity.check_parity(a, r2c.cpu().numpy().astype(np.int64), p.cpu().numpy(), ref, 1e-5 / cols) == "exact"
    b = oracle_lib.generate(8, 256, 2)
    b[3, :] = 0.0                # every rank sees the same replicated state: every rank reports the row
    with pytest.raises(ab.B200LsapError) as e:
        _run_fused(oracle_lib, 8, 256, 0, 2, a=b)
    assert e.value.status == ab.ERR_INFEASIBLE
    h = ab.Handle(0)
    t = torch.zeros((8, 64), dtype=torch.float32, device="cuda")
    out_t = torch.zeros(8, dtype=torch.int32, device="cuda")
    with pytest.raises(ab.B200LsapError):
        h.solve_sharded_device(t.data_ptr(), 8, 64, 64, 0, 64, out_t.data_ptr())     # mailboxes not connected
    h.shard_comm_create(4, 1)
    h.shard_comm_connect(0, 1)
    with pytest.raises(ab.B200LsapError):
        h.solve_sharded_device(t.data_ptr(), 8, 64, 64, 0, 64, out_t.data_ptr())     # more rows than rows_cap
    with pytest.raises(ab.B200LsapError):
        h.shard_comm_create(4, 17)                                                   # more than 16 ranks
    h.close()


def test_fused_sharded_missing_peer_times_out_instead_of_hanging(oracle_lib, monkeypatch):
    # rank 1 never starts: rank 0 gives up after the spin limit (50 ms here) with an error status
    import torch
    import auctionalgorithm_b200 as ab
    from auctionalgorithm_b200 import sharded
    monkeypatch.setenv("B200LSAP_SHARD_SPIN_MS", "50")
    handles = [ab.Handle(0) for _ in range(2)]
    sharded.connect_fused_local(handles, 64)
    t = torch.from_numpy(oracle_lib.generate(64, 64, 1)).cuda()
    out_t = torch.zeros(64, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    with pytest.raises(ab.B200LsapError) as e:
        handles[0].solve_sharded_device(t.data_ptr(), 64, 64, 64, 0, 128, out_t.data_ptr())
    assert e.value.status == ab.ERR_CUDA and "peer" in str(e.value)
    for h in handles:
        h.close()
