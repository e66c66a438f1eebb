"""Scratch timing of the device-resident paths (not the judged bench: see bench.py)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import auctionalgorithm_b200 as ab

h = ab.Handle(0)
def dense(rows, cols, seed=1, reps=3, **opt):
    t = torch.empty((rows, cols), dtype=torch.float32, device="cuda")
    h.generate_device(t.data_ptr(), 1, rows, cols, cols, 1, seed)
    r2c = torch.empty(rows, dtype=torch.int32, device="cuda")
    best = None
    for _ in range(reps):
        st = h.solve_device(t.data_ptr(), rows, cols, cols, 1, r2c.data_ptr(), **opt)
        if best is None or st["ms_solve"] < best["ms_solve"]: best = st
    gbs = best["bid_bytes"] / best["ms_solve"] / 1e6
    print(f"dense {rows}x{cols} {opt}: {best['ms_solve']:.3f} ms rounds {best['rounds']} bids {best['bids']} phases {best['phases']} narrow {best['narrow_rounds']} tail {best['tail_bids']}/{best['tail_runs']} "
          f"us/round {1e3*best['ms_solve']/max(best['rounds'],1):.2f} alg GB/s {gbs:.0f} "
          f"cache hit/miss {best['cache_hits']}/{best['cache_misses']} eps0 {best['eps_first']:.3g}", flush=True)
    dbg = h.debug_counters()
    if any(dbg):
        print("   dbg cycles:", dbg, flush=True)
    ms, nb = h.bid_sweep_device(t.data_ptr(), rows, cols, cols, True)
    ms2, nb = h.bid_sweep_device(t.data_ptr(), rows, cols, cols, True)
    print(f"   full bid sweep: {ms2:.4f} ms -> {nb/ms2/1e6:.0f} GB/s", flush=True)
    return r2c

def batch(n, rows=32, cols=48, reps=3):
    t = torch.empty((n, rows, cols), dtype=torch.float32, device="cuda")
    h.generate_device(t.data_ptr(), n, rows, cols, cols, 1, 1)
    out = torch.empty((n, rows), dtype=torch.int32, device="cuda")
    best = 1e9
    for _ in range(reps):
        st = h.solve_batch_device(t.data_ptr(), n, rows, cols, out.data_ptr())
        best = min(best, st["ms_solve"])
    print(f"batch {n}x{rows}x{cols}: {best:.3f} ms -> {n/best/1e3:.1f} M problems/s, {st['bid_bytes']/best/1e6:.0f} GB/s", flush=True)

def batch_shards(n=1 << 20, shards=8, reps=3, **opt):
    """C3 on the eight shards bench.py hands to ranks 0..7 (problem0 = k * n): a lucky shard cannot be the number."""
    t = torch.empty((n, 32, 48), dtype=torch.float32, device="cuda")
    out = torch.empty((n, 32), dtype=torch.int32, device="cuda")
    res = []
    for k in range(shards):
        h.generate_device(t.data_ptr(), n, 32, 48, 48, 1, 1, problem0=k * n)
        best = 1e9
        for _ in range(reps):
            st = h.solve_batch_device(t.data_ptr(), n, 32, 48, out.data_ptr(), **opt)
            best = min(best, st["ms_solve"])
        res.append(best)
        print(f"shard {k} {opt}: {best:.3f} ms -> {n/best/1e3:.1f} M problems/s, takeovers {st['batch_takeovers']}, "
              f"max bids {st['batch_max_bids']}, takeover hits/misses {st['cache_hits']}/{st['cache_misses']}", flush=True)
    print(f"shards {opt}: best {min(res):.3f} worst {max(res):.3f} ms (worst/best {max(res)/min(res):.3f})", flush=True)

if __name__ == "__main__":
    which = sys.argv[1:] or ["c1", "c2", "c3", "sq"]
    if "c3shards" in which:
        batch_shards()
    if "monster" in which:   # the slowest problems of the bench shards alone: takeover cost per bid
        for pid in (3097078, 1117884, 3248947):
            t = torch.empty((1, 32, 48), dtype=torch.float32, device="cuda")
            out = torch.empty((1, 32), dtype=torch.int32, device="cuda")
            h.generate_device(t.data_ptr(), 1, 32, 48, 48, 1, 1, problem0=pid)
            for cap in (96, -1):
                best = 1e9
                for _ in range(5):
                    st = h.solve_batch_device(t.data_ptr(), 1, 32, 48, out.data_ptr(), batch_bid_cap=cap)
                    best = min(best, st["ms_solve"])
                print(f"problem {pid} cap {cap}: {best*1e3:.1f} us, bids {st['batch_max_bids']}, hits/misses "
                      f"{st['cache_hits']}/{st['cache_misses']} -> {best*1e3/st['batch_max_bids']:.3f} us/bid", flush=True)
                dbg = h.debug_counters()
                if any(dbg[:4]):
                    print("   takeover cycles [eval, shfl, update, next-row]:", dbg[:4], flush=True)
    if "c5" in which:   # 65536 x 98304 on one GPU: the dense solver, its wide rounds alone, the fused sharded kernel, one sweep
        from auctionalgorithm_b200 import sharded
        rows, cols = 65536, 98304
        t = torch.empty((rows, cols), dtype=torch.float32, device="cuda")
        h.generate_shard_device(t.data_ptr(), rows, cols, cols, 0, cols, 1)
        r2c = torch.empty(rows, dtype=torch.int32, device="cuda")
        for opt in (dict(), dict(tail_rows=-1, narrow_rows=-1)):
            for _ in range(2):
                st = h.solve_device(t.data_ptr(), rows, cols, cols, 1, r2c.data_ptr(), **opt)
            print(f"c5 dense {opt}: {st['ms_solve']:.3f} ms rounds {st['rounds']} bids {st['bids']} narrow {st['narrow_rounds']} launches {st['launches']}", flush=True)
        for _ in range(3):
            ms, nb = h.bid_sweep_device(t.data_ptr(), rows, cols, cols, False)
        print(f"c5 first sweep: {ms:.3f} ms -> {nb/ms/1e6:.0f} GB/s", flush=True)
        h.shard_comm_create(rows, 1); h.shard_comm_connect(0, 1)
        for _ in range(2):
            r, p, st = sharded.solve_fused(h, t, rows, cols, cols, 0, cols)
        print(f"c5 fused N=1: {st['ms_solve']:.3f} ms rounds {st['rounds']} bids {st['bids']}", flush=True)
        dbg = h.debug_counters()
        if any(dbg):
            print("   dbg:", dbg, flush=True)
    if "c5n" in which:   # bidders per round of C5 (host-driven form, one shard) with the kernels' times
        rows, cols = 65536, 98304
        t = torch.empty((rows, cols), dtype=torch.float32, device="cuda")
        h.generate_shard_device(t.data_ptr(), rows, cols, cols, 0, cols, 1)
        amax = h.shard_begin(t.data_ptr(), rows, cols, cols, 0, cols)
        h.shard_set_amax(amax)
        send = torch.empty(rows * 32, dtype=torch.uint8, device="cuda")
        n, out = rows, []
        while n > 0:
            torch.cuda.synchronize(); t0 = time.perf_counter()
            h.shard_bid(send.data_ptr()); torch.cuda.synchronize(); t1 = time.perf_counter()
            n2 = h.shard_resolve(send.data_ptr(), 1); t2 = time.perf_counter()
            out.append((n, (t1 - t0) * 1e6, (t2 - t1) * 1e6)); n = n2
        print("c5 rounds (bidders, bid us, resolve us):", [(a, round(b), round(c)) for a, b, c in out], flush=True)
    if "monster1" in which:  # one launch of the slowest problem alone (ncu source-level profile of the takeover)
        t = torch.empty((1, 32, 48), dtype=torch.float32, device="cuda")
        out = torch.empty((1, 32), dtype=torch.int32, device="cuda")
        h.generate_device(t.data_ptr(), 1, 32, 48, 48, 1, 1, problem0=3097078)
        for _ in range(3):
            st = h.solve_batch_device(t.data_ptr(), 1, 32, 48, out.data_ptr())
        print(f"monster alone: {st['ms_solve']*1e3:.1f} us", flush=True)
    if "c3host" in which:   # the host entry of the batched mode: pinned input, pageable and pinned output
        n = 1 << 18
        t = torch.empty((n, 32, 48), dtype=torch.float32, device="cuda")
        h.generate_device(t.data_ptr(), n, 32, 48, 48, 1, 1)
        host = torch.empty((n, 32, 48), dtype=torch.float32, pin_memory=True); host.copy_(t); torch.cuda.synchronize()
        a = host.numpy()
        for _ in range(2):
            h.solve_batch(a)
        for rep in range(3):
            t0 = time.perf_counter(); r = h.solve_batch(a); sec = time.perf_counter() - t0
            print(f"c3 host entry: {n/sec/1e6:.2f} M problems/s, {a.nbytes/sec/1e9:.1f} GB/s H2D, {sec*1e3:.2f} ms", flush=True)
    if "c3s3" in which:     # three shards (0, 1, 2) only: A/B runs with environment knobs
        batch_shards(shards=3)
    if "c3caps" in which:
        for cap in (-1, 64, 96, 128, 192):
            batch_shards(shards=3, batch_bid_cap=cap)
    if "c1" in which: dense(200, 300)
    if "c2" in which: dense(2048, 3072)
    if "c3" in which: batch(1 << 16); batch(1 << 20)
    if "c3s" in which: batch(1 << 18)
    if "sweep" in which:
        t = torch.empty((16384, 16384), dtype=torch.float32, device="cuda")
        h.generate_device(t.data_ptr(), 1, 16384, 16384, 16384, 1, 1)
        for _ in range(4):
            ms, nb = h.bid_sweep_device(t.data_ptr(), 16384, 16384, 16384, False)
            print(f"sweep 16384^2: {ms:.4f} ms -> {nb/ms/1e6:.0f} GB/s", flush=True)
    if "sweep2" in which:  # zero prices and the prices a solve leaves; min of 8 launches each
        t = torch.empty((16384, 16384), dtype=torch.float32, device="cuda")
        h.generate_device(t.data_ptr(), 1, 16384, 16384, 16384, 1, 1)
        a = [h.bid_sweep_device(t.data_ptr(), 16384, 16384, 16384, False)[0] for _ in range(8)]
        r2c = torch.empty(16384, dtype=torch.int32, device="cuda")
        st = h.solve_device(t.data_ptr(), 16384, 16384, 16384, 1, r2c.data_ptr())
        st = h.solve_device(t.data_ptr(), 16384, 16384, 16384, 1, r2c.data_ptr())
        b = [h.bid_sweep_device(t.data_ptr(), 16384, 16384, 16384, True)[0] for _ in range(8)]
        nb = 16384 * 16384 * 4
        print(f"sweep2: zero prices min {min(a):.4f} mean {sum(a[3:])/5:.4f} ms -> {nb/min(a)/1e6:.0f} GB/s | last prices min {min(b):.4f} "
              f"mean {sum(b[3:])/5:.4f} ms -> {nb/(sum(b[3:])/5)/1e6:.0f} GB/s | solve {st['ms_solve']:.2f} ms", flush=True)
    if "sparse" in which or "sparse20k" in which:
        import scipy.sparse as sp
        for n, per_row in (((20000, 40),) if "sparse20k" in which else ((5000, 5000), (20000, 40), (200000, 32), (1000000, 16))):
            rng = np.random.default_rng(1)
            if per_row >= n:
                a = (rng.integers(0, 1 << 24, (n, n)) + 1).astype(np.float64) / float(1 << 24)
                m = sp.csr_matrix(a)
            else:
                ci = rng.integers(0, n, (n, per_row)); ci[:, 0] = rng.permutation(n); ci.sort(axis=1)
                keep = np.ones_like(ci, bool); keep[:, 1:] = ci[:, 1:] != ci[:, :-1]
                rp = np.concatenate([[0], np.cumsum(keep.sum(axis=1))]).astype(np.int32)
                vals = (rng.integers(0, 1 << 24, int(keep.sum())) + 1) / float(1 << 24)
                m = sp.csr_matrix((vals, ci[keep].astype(np.int32), rp), shape=(n, n))
            for dt in (np.float64, np.float32):
                v = m.data.astype(dt)
                for _ in range(2):
                    t0 = time.time(); r2c = h.solve_csr(v, m.indptr, m.indices, n, n); wall = time.time() - t0
                st = h.last_stats
                print(f"sparse {n}x{n} nnz {m.nnz} {dt.__name__}: solve {st['ms_solve']:.2f} ms (upload {st['ms_upload']:.1f}, wall {wall*1e3:.1f}) "
                      f"rounds {st['rounds']} narrow {st['narrow_rounds']} bids {st['bids']} phases {st['phases']} "
                      f"objective {m[np.arange(n), r2c].sum():.6f}", flush=True)
    if "sparse_eps" in which:
        for n, per_row in ((20000, 40), (200000, 32)):
            rng = np.random.default_rng(1)
            ci = rng.integers(0, n, (n, per_row)); ci[:, 0] = rng.permutation(n); ci.sort(axis=1)
            keep = np.ones_like(ci, bool); keep[:, 1:] = ci[:, 1:] != ci[:, :-1]
            rp = np.concatenate([[0], np.cumsum(keep.sum(axis=1))]).astype(np.int32)
            vals = ((rng.integers(0, 1 << 24, int(keep.sum())) + 1) / float(1 << 24)).astype(np.float32)
            cidx = ci[keep].astype(np.int32)
            for e0 in (0.25, 0.05, 0.01, 0.002, 0.0004):
                for f in (4.0, 7.0, 16.0):
                    h.solve_csr(vals, rp, cidx, n, n, eps_start=e0, eps_factor=f)
                    st = h.last_stats
                    print(f"sparse {n} eps0 {e0} factor {f}: {st['ms_solve']:.1f} ms rounds {st['rounds']} bids {st['bids']} phases {st['phases']}", flush=True)
    if "sq" in which:
        dense(2048, 2048); dense(4096, 4096)
    if "c4f" in which:
        for f in (3.0, 4.0, 5.0, 7.0, 10.0):
            dense(16384, 16384, reps=1, eps_factor=f)
    if "seeds" in which:
        for f in (5.0, 7.0):
            for n in (4096, 8192, 16384):
                for seed in (1, 2, 3):
                    dense(n, n, seed=seed, reps=1, eps_factor=f)
    if "struct" in which:
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
        import test_gpu_structured as tg
        for gen in tg.GENERATORS:
            a = np.ascontiguousarray(gen(np.random.default_rng(7), 2048, 2048))
            for opt in (dict(), dict(eps_start=0.25), dict(row_cache=-1, eps_start=0.25)):
                t0 = time.time()
                try:
                    r2c, _ = h.solve(a, want_prices=True, max_rounds=3000000, **opt)
                    st = h.last_stats
                    print(f"{gen.__name__:16s} {str(opt):42s}: {st['ms_solve']:9.2f} ms rounds {st['rounds']:8d} phases {st['phases']:2d} "
                          f"hit/miss {st['cache_hits']}/{st['cache_misses']} eps0 {st['eps_first']:.3g} restarts {st['tail_runs']}", flush=True)
                except Exception as e:
                    print(f"{gen.__name__:16s} {str(opt):42s}: FAILED after {time.time()-t0:.1f} s: {e}", flush=True)
    if "lsap" in which:
        # the reference driver's own case (LSAP.cpp:23-33): 5000 x 5000 double, (Random + 1) / max
        rng = np.random.default_rng(1)
        a = rng.uniform(-1.0, 1.0, (5000, 5000)) + 1.0
        a /= a.max()
        a[a == 0.0] = 1e-12
        for opt in (dict(), dict(row_cache=-1, eps_start=0.25), dict(eps_scaling=False)):
            if opt.get("eps_scaling") is False and "slow" not in which:
                continue
            r2c, _ = h.solve(np.asfortranarray(a), want_prices=True, **opt)
            st = h.last_stats
            print(f"LSAP.cpp case 5000x5000 double {opt}: solve {st['ms_solve']:.2f} ms (upload {st['ms_upload']:.1f}) rounds {st['rounds']} "
                  f"phases {st['phases']} hit/miss {st['cache_hits']}/{st['cache_misses']} objective {a[np.arange(5000), r2c].sum():.6f}", flush=True)
        af = a.astype(np.float32)
        r2c, _ = h.solve(np.asfortranarray(af), want_prices=True)
        st = h.last_stats
        print(f"same in float: solve {st['ms_solve']:.2f} ms rounds {st['rounds']} objective {af[np.arange(5000), r2c].astype(np.float64).sum():.6f}", flush=True)
    if "c4t" in which:
        dense(16384, 16384, reps=2)
    if "c4n" in which:
        dense(16384, 16384, reps=2)
        dense(16384, 16384, reps=1, eps_start=0.25)
        dense(16384, 16384, reps=1, row_cache=-1)
        dense(16384, 16384, reps=1, row_cache=-1, eps_start=0.25)
        dense(16384, 16384, reps=1, seed=2)
        dense(16384, 16384, reps=1, seed=3)
        dense(8192, 8192, reps=1)
    if "c4" in which:
        for opt in (dict(), dict(tail_rows=1), dict(tail_rows=4), dict(tail_rows=8), dict(narrow_rows=-1, tail_rows=8),
                    dict(narrow_rows=-1, tail_rows=-1)):
            dense(16384, 16384, reps=2, **opt)
    if "c4d" in which:
        dense(16384, 16384, reps=1)
        dense(16384, 16384, reps=1, tail_rows=-1)
        dense(16384, 16384, reps=1, tail_rows=1)
    if "c4" in which: dense(16384, 16384, reps=1)
