"""One gloo rank of tests/test_sharded_cpu.py: runs auctionalgorithm_b200.sharded.solve_sharded with the
numpy stand-in engine and writes its replica's result."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch.distributed as dist

from auctionalgorithm_b200 import sharded
from oracle import pyoracle as po
from shard_ref import NumpyShardEngine


def main():
    rows, cols, seed, out = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    col0, cl = sharded.split_columns(cols, world, align=4)[rank]
    a = po.generate(rows, cols, seed)[:, col0:col0 + cl]          # this rank only ever sees its shard
    r2c, prices, stats = sharded.solve_sharded(NumpyShardEngine(a, col0, cols), rows, want_prices=True)
    np.savez(f"{out}.{rank}.npz", r2c=r2c.numpy(), prices=prices.numpy(), rounds=stats["rounds"], bids=stats["bids"],
             exchanged=stats["exchanged_bytes"])
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
