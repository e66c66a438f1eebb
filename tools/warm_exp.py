import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
import auctionalgorithm_b200 as ab
from oracle import pyoracle as po
h = ab.Handle(0)
for n, frac, sd in ((512, 0.02, 0.01), (2048, 0.02, 0.01), (2048, 0.001, 0.01), (2048, 0.02, 0.001)):
    rng = np.random.default_rng(3)
    a1 = po.generate(n, n, 5)
    r1, p1 = h.solve(a1, want_prices=True)
    a2 = a1.copy(); idx = rng.random(a1.shape) < frac
    a2[idx] = np.clip(a2[idx] + rng.normal(0, sd, idx.sum()).astype(np.float32), 2.0**-24, 1.0).astype(np.float32)
    cold, pc = h.solve(a2, want_prices=True); c = dict(h.last_stats)
    res = [f"n {n} frac {frac} sd {sd}: cold rounds {c['rounds']} {c['ms_solve']:.2f} ms"]
    for name, opt in (("warm eps_final only", dict(eps_scaling=0)), ("warm auto scaling", dict(eps_scaling=1)),
                      ("warm start 1e-3", dict(eps_scaling=1, eps_start=1e-3)), ("warm start 1e-4", dict(eps_scaling=1, eps_start=1e-4)),
                      ("warm start 1e-5", dict(eps_scaling=1, eps_start=1e-5))):
        w, pw = h.solve(a2, want_prices=True, prices_in=p1, **opt); s = h.last_stats
        res.append(f"   {name}: rounds {s['rounds']} phases {s['phases']} {s['ms_solve']:.2f} ms same {np.array_equal(w, cold)} eps0 {s['eps_first']:.2g}")
    print("\n".join(res), flush=True)
