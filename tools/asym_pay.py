"""Scratch: rounds of the fixed-eps and the eps-scaled solve on near-square structured inputs."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import auctionalgorithm_b200 as ab
import test_gpu_structured as tg
h = ab.Handle(0)
for rows, cols in ((2048, 2052), (4096, 4100))[:int(os.environ.get("ASYM_SIZES", "2"))]:
    for gen in [g for g in tg.GENERATORS if g.__name__ not in os.environ.get("ASYM_SKIP", "").split(",")] + [None]:
        rng = np.random.default_rng(11)
        a = np.ascontiguousarray(gen(rng, rows, cols)) if gen else ((rng.integers(0, 1 << 24, (rows, cols)) + 1) / float(1 << 24)).astype(np.float32)
        name = gen.__name__ if gen else "uniform"
        out = []
        for opt in (dict(), dict(eps_scaling=1), dict(eps_scaling=1, eps_start=0.25))[:int(os.environ.get("ASYM_OPTS", "3"))]:
            t0 = time.time()
            try:
                h.solve(a, max_rounds=3_000_000, **opt); s = h.last_stats
                out.append(f"{s['rounds']} r / {s['ms_solve']:.1f} ms / rep {s['warm_fallback']}")
            except Exception as e:
                out.append(f"ERR {str(e)[:40]} {time.time()-t0:.1f}s")
        print(rows, cols, name, " | ".join(out), flush=True)
