"""Scratch: eps scaling on rows < cols, one case per line with rounds / repairs / time (run under `timeout`)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import auctionalgorithm_b200 as ab
from oracle import pyoracle as po
h = ab.Handle(0)
cases = [(3, 7, 2), (8, 12, 1), (32, 48, 3), (100, 101, 1), (200, 300, 1), (511, 513, 1), (512, 768, 1), (1024, 1536, 1), (2048, 3072, 2)]
for rows, cols, seed in cases:
    a = po.generate(rows, cols, seed)
    for opt in (dict(eps_scaling=1), dict(eps_scaling=1, eps_start=0.25), dict(eps_scaling=1, row_cache=-1)):
        t0 = time.time()
        try:
            r = h.solve(a, max_rounds=2_000_000, **opt)
            s = h.last_stats
            print(rows, cols, opt, "rounds", s["rounds"], "phases", s["phases"], "repairs", s["warm_fallback"], f"{time.time()-t0:.2f}s", flush=True)
        except Exception as e:
            print(rows, cols, opt, "ERR", e, f"{time.time()-t0:.2f}s", flush=True)
