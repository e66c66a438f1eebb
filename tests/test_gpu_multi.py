"""Multi-GPU tests (skipped with fewer than two GPUs): the column-sharded solve on REAL ranks, one process per GPU
launched by torch.distributed.run, against reference-made golden vectors.  Run with `gpurun --gpus 2`."""
import json
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    import torch
    return torch.cuda.device_count()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("name,world", [("auction_8192x12288_s1", 2), ("auction_2048x3072_s1", 2), ("auction_2048x2048_s1", 2),
                                        ("auction_8192x12288_s1", 4), ("auction_8192x12288_s1", 8)])
def test_sharded_solve_on_real_ranks_matches_the_reference(tmp_path, name, world):
    if _gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    out = str(tmp_path / "res.json")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "_fused_worker.py"), name, out]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    res = json.load(open(out))
    assert res["fused_exact"], res
    assert res["replicas_identical"] and res["host_driven_equal"], res
    assert res["eps_cs_violation_local_columns"] <= res["eps"] * (1 + 1e-9) + 1e-15, res   # a few ulps of the fp64 check itself


def test_multi_handle_solves_across_devices_in_one_process(oracle_lib):
    """b200lsap_solve_sharded_* from ONE host thread over two devices (peer access, mailboxes wired by pointer)."""
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    import numpy as np
    import torch
    import auctionalgorithm_b200 as ab
    from auctionalgorithm_b200 import sharded
    import parity
    g = np.load(f"{parity.GOLDEN_DIR}/auction_2048x3072_s1.npz")
    rows, cols, seed = 2048, 3072, 1
    handles = [ab.Handle(0), ab.Handle(1)]
    sharded.connect_fused_local(handles, rows, devices=[0, 1])
    parts = sharded.split_columns(cols, 2)
    a = oracle_lib.generate(rows, cols, seed)
    shards = [torch.from_numpy(np.ascontiguousarray(a[:, c0:c0 + cl])).to(f"cuda:{d}") for d, (c0, cl) in enumerate(parts)]
    res = sharded.solve_fused_local(handles, shards, rows, parts, [cl for _, cl in parts], cols, want_prices=True)
    for r2c, p, st in res:
        assert np.array_equal(r2c.cpu().numpy().astype(np.int64), parity.golden_row_to_col(g))
    assert torch.equal(res[0][1].cpu(), res[1][1].cpu())
    for h in handles:
        h.close()
