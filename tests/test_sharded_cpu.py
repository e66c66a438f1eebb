"""CPU tests of the N > 1 path: the column split and the sharded host loop under gloo, world_size 2
and 3, with the numpy stand-in engine (tests/shard_ref.py) in the place of the CUDA kernels."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import parity
from auctionalgorithm_b200 import sharded

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("cols,n", [(300, 2), (98304, 8), (98304, 4), (1000, 3), (129, 2), (16, 4), (5000, 7)])
def test_split_columns_covers_every_column_once(cols, n):
    parts = sharded.split_columns(cols, n)
    assert len(parts) == n and parts[0][0] == 0
    for (c0, cl), (d0, _) in zip(parts, parts[1:]):
        assert c0 + cl == d0
    assert parts[-1][0] + parts[-1][1] == cols
    assert all(c0 % 4 == 0 and cl > 0 for c0, cl in parts)
    sizes = [cl for _, cl in parts]
    assert max(sizes) - min(sizes) <= 128 + 127


def test_split_columns_rejects_impossible_splits():
    with pytest.raises(ValueError):
        sharded.split_columns(3, 4)
    with pytest.raises(ValueError):
        sharded.split_columns(8, 3)   # multiples of 4 only


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("rows,cols,seed,world", [(60, 90, 5, 2), (48, 48, 6, 2), (40, 100, 7, 3)])
def test_sharded_loop_under_gloo_matches_the_oracle(tmp_path, oracle_lib, rows, cols, seed, world):
    out = str(tmp_path / "res")
    port = _free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "_sharded_worker.py"), str(rows),
                                       str(cols), str(seed), out], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT))
    for p in procs:
        log, _ = p.communicate(timeout=300)
        assert p.returncode == 0, log.decode()[-2000:]
    res = [np.load(f"{out}.{r}.npz") for r in range(world)]
    for r in res[1:]:   # replicas end identical, price for price
        assert np.array_equal(r["r2c"], res[0]["r2c"]) and np.array_equal(r["prices"], res[0]["prices"])
        assert int(r["rounds"]) == int(res[0]["rounds"])
    a = oracle_lib.generate(rows, cols, seed)
    ref_out = oracle_lib.oracle_solve(a.astype(np.float64))
    ref = np.full(rows, -1, np.int64)
    ref[ref_out["x"]] = ref_out["y"]
    parity.check_parity(a, res[0]["r2c"].astype(np.int64), res[0]["prices"], ref, 1e-5 / cols)
    assert int(res[0]["exchanged"]) == world * 32 * int(res[0]["bids"])
