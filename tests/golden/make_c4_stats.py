"""Runs the C oracle (bit-identical to the reference, tests/test_oracle.py) on the C4 input
(16384^2, seed 1) to record the work counters the unmodified reference cannot report:
bids, sweeps, evictions.  bench.py's cpu_baseline extrapolates a bounded sample with them.
~20 min of one core.  Writes tests/golden/c4_oracle_stats.json."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po
n = 16384
a = np.asfortranarray(po.generate(n, n, 1).astype(np.float64))
t = time.time()
o = po.oracle_solve(a)
sec = time.time() - t
g = np.load(os.path.join(ROOT, "tests/golden/auction_16384x16384_s1.npz"))
same = bool(np.array_equal(o["x"], g["x"]) and np.array_equal(o["y"], g["y"]))
out = dict(rows=n, cols=n, seed=1, seconds_oracle=sec, seconds_reference=float(g["seconds"]),
           identical_to_reference_edges=same, objective=float(o["v"].sum()), **o["stats"])
json.dump(out, open(os.path.join(ROOT, "tests/golden/c4_oracle_stats.json"), "w"), indent=1)
print(out)
