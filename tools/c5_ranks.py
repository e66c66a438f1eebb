"""Scratch: the C5 leg of bench.py alone under torchrun (one line per run), for A/B of the fused sharded kernel."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import auctionalgorithm_b200 as ab
from auctionalgorithm_b200 import sharded

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
h = ab.Handle(local)
rows, cols = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (65536, 98304)
res = sharded.bench_leg(h, torch, dist, rank, world, 3, rows, cols, 1)
if rank == 0:
    print("c5", world, "ranks:", json.dumps({k: res[k] for k in ("ms_per_solve", "rounds", "us_per_round", "parity", "eps_cs_holds")}), flush=True)
if world > 1:
    dist.destroy_process_group()
