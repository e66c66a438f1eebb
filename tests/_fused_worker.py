"""One NCCL rank of tests/test_gpu_multi.py (launched by torch.distributed.run, one process per GPU): solves a
reference-made golden problem column-sharded over the ranks, with the fused kernel (in-kernel exchange over peer
memory) and with the host-driven form (one NCCL all-gather per round), and writes what rank 0 saw."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import auctionalgorithm_b200 as ab  # noqa: E402
from auctionalgorithm_b200 import sharded  # noqa: E402
import parity  # noqa: E402

name, out_path = sys.argv[1], sys.argv[2]
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
g = np.load(os.path.join(parity.GOLDEN_DIR, name + ".npz"))
rows, cols, seed = int(g["rows"]), int(g["cols"]), int(g["seed"])
c0, cl = sharded.split_columns(cols, world)[rank]
h = ab.Handle(local)
t = torch.empty((rows, cl), dtype=torch.float32, device="cuda")
h.generate_shard_device(t.data_ptr(), rows, cl, cl, c0, cols, seed)
torch.cuda.synchronize()
# fused: twice on the same wiring
sharded.connect_fused(h, rows)
for _ in range(2):
    r2c, prices, st = sharded.solve_fused(h, t, rows, cl, cl, c0, cols, want_prices=True)
ref = torch.from_numpy(parity.golden_row_to_col(g)).cuda()
res = {"fused_exact": bool(torch.equal(r2c.long(), ref)), "fused_rounds": st["rounds"], "fused_ms": st["ms_solve"],
       "world": world}
# every replica identical
mine = torch.cat([r2c.double(), prices])
allr = [torch.empty_like(mine) for _ in range(world)]
dist.all_gather(allr, mine)
res["replicas_identical"] = all(bool(torch.equal(x, allr[0])) for x in allr)
# host-driven form over NCCL: same rounds, same prices
eng = sharded.CudaShardEngine(h, t, rows, cl, cl, c0, cols)
r2, p2, st2 = sharded.solve_sharded(eng, rows, want_prices=True)
res["host_driven_equal"] = bool(torch.equal(r2, r2c) and torch.equal(p2, prices) and st2["rounds"] == st["rounds"])
res["eps_cs_violation_local_columns"] = sharded.eps_cs_violation(t, r2c, prices, c0)
res["eps"] = 1e-5 / cols
if rank == 0:
    json.dump(res, open(out_path, "w"))
h.close()
dist.destroy_process_group()
