#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 auction path (contract: task brief §4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c4|c2|c1]

Default workload: BASELINE.json configs[3] = C4, the configuration the metric is quoted on: one dense
16384 x 16384 fp32 problem (1 GiB, larger than L2) solved to the reference's assignment.  A step = one
full solve.
  value         solve ms per problem, matrix already resident in HBM (CUDA events on the stream)
  e2e           the same solve through the reference-facing host API (Eigen-layout column-major pinned
                host matrix in, row->column out): H2D + transpose + solve + D2H inside the timed region
  roofline      the bidding kernel alone (every row bids once, one launch), timed live with CUDA events
  cpu_baseline  the CPU port of Auction<double>::solve (oracle/, pinned bit-for-bit against the compiled
                reference) on a bounded sample of the same input, scaled by the reference's bid count
  mht           BASELINE's second quantity: batched 32x48 problems/s (C3), sharded over the N ranks
  sharded       BASELINE configs[4]: 65536 x 98304 column-sharded over the N ranks, one all-gather of
                32-byte records per round (N = 1: the same code path with a single shard)
N > 1: a single mid-size dense problem does not shard ("replicas only"): every rank solves
its own replica and `value` is the aggregate ms per solve (step time / N); `mht` and `sharded` are the
paths that shard.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c1": dict(rows=200, cols=300), "c2": dict(rows=2048, cols=3072), "c4": dict(rows=16384, cols=16384),
}
C5 = dict(rows=65536, cols=98304)
SEED = 1


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def peaks() -> tuple[float, str]:
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def ncu_traffic():
    """DRAM bytes per launch of the bidding kernel from the committed ncu --set full summary."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "kbid_ncu_summary.json")))["dram_bytes_per_launch"]
    except Exception:
        return None


def c4_reference_counters() -> dict:
    p = os.path.join(ROOT, "tests", "golden", "c4_oracle_stats.json")
    return json.load(open(p)) if os.path.exists(p) else {}


class CpuArm:
    """The reference's CPU path on the host cores.  Auction<Scalar>::solve is single-threaded by
    construction (Auction.h:44-109); AuctionMT is timed beside it on a small case because it is slower
    than the single thread at every size (two barrier crossings per bid)."""

    def __init__(self, rows: int, cols: int):
        import numpy as np
        from oracle import pyoracle as po
        po.build()
        self.po, self.np, self.rows, self.cols = po, np, rows, cols
        self.a = np.asfortranarray(po.generate(rows, cols, SEED).astype(np.float64))  # Eigen layout, widened fp32
        self.full = rows * cols <= 4096 * 6144 and rows != cols
        self.budget = None

    def sample(self, seconds_target: float) -> dict:
        po, a, rows, cols = self.po, self.a, self.rows, self.cols
        if self.full:
            if po.reference() is not None:
                out = po.reference_solve(a, "auction")
                return dict(value=out["seconds"] * 1e3, unit="ms", cores=1, kind="reference",
                            sample=f"full Auction<double>::solve of {rows}x{cols} seed {SEED} (unmodified reference headers)")
            t = time.perf_counter()
            out = po.oracle_solve(a)
            sec = time.perf_counter() - t
            return dict(value=sec * 1e3, unit="ms", cores=1, kind="port",
                        sample=f"full Auction<double> solve of {rows}x{cols} seed {SEED}: {out['stats']['bids']} bids")
        if self.budget is None:
            po.oracle_solve(a, max_bids=1000)
            t = time.perf_counter()
            probe = po.oracle_solve(a, max_bids=3000)
            per_bid = (time.perf_counter() - t) / max(probe["stats"]["bids"], 1)
            self.budget = max(3000, int(seconds_target / per_bid))
        t = time.perf_counter()
        out = po.oracle_solve(a, max_bids=self.budget)
        sec = time.perf_counter() - t
        bids = out["stats"]["bids"]
        if out["status"] == po.ORC_OK:
            return dict(value=sec * 1e3, unit="ms", cores=1, kind="port", sample=f"full solve, {bids} bids")
        ref = c4_reference_counters() if (rows, cols) == (16384, 16384) else {}
        if ref.get("bids"):
            total = ref["bids"]
            note = (f"first {bids} bids of Auction<double> (CPU port, bit-identical to the reference) on {rows}x{cols} "
                    f"seed {SEED} took {sec:.1f} s ({1e6 * sec / bids:.0f} us/bid), scaled to the {total} bids of the "
                    f"full reference solve; the unmodified reference measured offline took "
                    f"{ref.get('seconds_reference', 0):.0f} s (tests/golden/c4_oracle_stats.json)")
            return dict(value=sec / bids * total * 1e3, unit="ms", cores=1, kind="port", sample=note)
        return dict(value=None, unit="ms", cores=1, kind="port",
                    sample=f"first {bids} bids took {sec:.1f} s ({1e6 * sec / bids:.0f} us/bid); total bid count unknown")

    def mt_note(self) -> dict:
        """AuctionMT (ThreadBarrier) beside Auction on a bounded square case, unmodified reference."""
        po, np = self.po, self.np
        if po.reference() is None:
            return {"unavailable": "oracle/_ref not built"}
        n = 1024
        a = np.asfortranarray(po.generate(n, n, SEED).astype(np.float64))
        st = po.reference_solve(a, "auction")["seconds"]
        cores = os.cpu_count() or 1
        out = {"case": f"{n}x{n} double, seed {SEED}", "auction_st_ms": st * 1e3, "host_cores": cores}
        for t in sorted({2, min(4, cores)}):
            out[f"auctionmt_{t}threads_ms"] = po.reference_solve(a, "mt", n_threads=t)["seconds"] * 1e3
        return out


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    w = WORKLOADS[args.workload]
    arm = CpuArm(w["rows"], w["cols"])
    vals, last = [], None
    for k in range(args.warmup + args.steps):
        last = arm.sample(seconds_target=6.0)
        if k >= args.warmup and last["value"] is not None:
            vals.append(last["value"])
    v = statistics.mean(vals) if vals else None
    last = dict(last, value=v)
    print(json.dumps({
        "impl": "reference", "metric": f"solve_ms_{w['rows']}x{w['cols']}_fp32", "value": v, "unit": "ms",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": v,
        "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}: dense {w['rows']}x{w['cols']} fp32 widened to f64, seed {SEED}, "
                               "the reference's Auction<double>::solve on the host, single thread (the solver is "
                               "single-threaded; AuctionMT is slower at every size, see cpu_baseline.auctionmt)"},
        "cpu_baseline": dict(last, auctionmt=arm.mt_note()),
        "e2e": {"value": v, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_sparse(handle, n: int = 20000, per_row: int = 40):
    """b200lsap_solve_csr_f32 on a seeded n x n matrix with about per_row stored entries per row (host arrays in,
    assignment out).  The unmodified reference's Auction<double, SparseMatrix<double, RowMajor>>::solve took 441 s
    on this input in the build container; it is not re-timed here."""
    import numpy as np
    rng = np.random.default_rng(1)
    ci = rng.integers(0, n, (n, per_row))
    ci[:, 0] = rng.permutation(n)              # a perfect matching exists
    ci.sort(axis=1)
    keep = np.ones_like(ci, bool)
    keep[:, 1:] = ci[:, 1:] != ci[:, :-1]      # one entry per (row, column)
    row_ptr = np.concatenate([[0], np.cumsum(keep.sum(axis=1))]).astype(np.int32)
    col_idx = ci[keep].astype(np.int32)
    values = ((rng.integers(0, 1 << 24, len(col_idx)) + 1) / float(1 << 24)).astype(np.float32)
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        r2c = handle.solve_csr(values, row_ptr, col_idx, n, n)
        wall = (time.perf_counter() - t0) * 1e3
        st = dict(handle.last_stats)
        st["wall_ms"] = wall
        if best is None or st["ms_solve"] < best["ms_solve"]:
            best = st
    ok = bool(len(np.unique(r2c)) == n)
    return {"workload": f"{n}x{n} fp32, {len(values)} stored entries (about {per_row} per row), numpy seed 1",
            "ms_solve": best["ms_solve"], "ms_e2e_host_arrays": best["wall_ms"], "rounds": best["rounds"],
            "narrow_rounds": best["narrow_rounds"], "bids": best["bids"], "phases": best["phases"],
            "us_per_round": 1e3 * best["ms_solve"] / max(best["rounds"], 1), "one_to_one": ok,
            "reference_cpu_s_offline": 441.0}


def run_mht(handle, torch, n_problems: int, rank: int, steps: int, warmup: int, with_host_legs: bool = False):
    """C3: this rank's shard of the batch, generated on the device (inputs resident in HBM)."""
    t = torch.empty((n_problems, 32, 48), dtype=torch.float32, device="cuda")
    out = torch.empty((n_problems, 32), dtype=torch.int32, device="cuda")
    handle.generate_device(t.data_ptr(), n_problems, 32, 48, 48, 1, SEED, problem0=rank * n_problems)
    for _ in range(warmup):
        handle.solve_batch_device(t.data_ptr(), n_problems, 32, 48, out.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        st = handle.solve_batch_device(t.data_ptr(), n_problems, 32, 48, out.data_ptr())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    extra = {}
    if with_host_legs:
        # end to end through the reference-facing batched entry (b200lsap_solve_batch_f32): HOST arrays in and out,
        # the library's own copies inside the timed call
        n_e2e = min(n_problems, 1 << 18)
        host_t = torch.empty((n_e2e, 32, 48), dtype=torch.float32, pin_memory=True)
        host_t.copy_(t[:n_e2e])
        torch.cuda.synchronize()
        host = host_t.numpy()
        handle.solve_batch(host[:1024])
        handle.solve_batch(host)
        t0 = time.perf_counter()
        r2c = handle.solve_batch(host)
        sec = time.perf_counter() - t0
        assert (r2c >= 0).all()
        extra["e2e"] = {"problems_per_s": n_e2e / sec, "problems": n_e2e, "h2d_bytes": int(host.nbytes),
                        "d2h_bytes": int(r2c.nbytes), "ms": sec * 1e3, "h2d_gbs": host.nbytes / sec / 1e9,
                        "input": "pinned host array [batch, 32, 48] fp32, pageable result array; bound by the "
                                 "6144 bytes per problem that cross PCIe, not by the kernel"}
        # the reference has no batched entry: Auction<float, MatrixXf>::solve in a loop, here the CPU
        # port of it (bit-identical, tests/test_oracle.py), one thread and all host threads, on a subsample
        from oracle import pyoracle as po
        cores = os.cpu_count() or 1
        n_cpu = min(n_e2e, 1 << 16)
        cpu = {"kind": "port", "cores": cores, "sample": f"first {n_cpu} problems of the same batch"}
        for key, dbl, thr, cnt in (("float_1thread", False, 1, n_cpu // 8), ("float_all_threads", False, cores, n_cpu),
                                   ("double_all_threads", True, cores, n_cpu)):
            t0 = time.perf_counter()
            ref, failed = po.oracle_batch(host[:cnt], use_double=dbl, n_threads=thr)
            cpu[key + "_problems_per_s"] = cnt / (time.perf_counter() - t0)
            if dbl:
                cpu["gpu_result_identical_to_double_port"] = bool(failed == 0 and (ref == r2c[:cnt]).all())
        extra["cpu_baseline"] = cpu
    del t, out
    return ms, st["bid_bytes"], extra


def run_sharded(handle, torch, dist, rank: int, world: int, steps: int, rows: int, cols: int):
    """C5: one rows x cols problem column-sharded over the ranks, generated per shard on the device."""
    from auctionalgorithm_b200 import sharded
    c0, cl = sharded.split_columns(cols, world)[rank]
    t = torch.empty((rows, cl), dtype=torch.float32, device="cuda")
    handle.generate_shard_device(t.data_ptr(), rows, cl, cl, c0, cols, SEED)
    eng = sharded.CudaShardEngine(handle, t, rows, cl, cl, c0, cols)

    def once():
        if world > 1:
            return sharded.solve_sharded(eng, rows)
        return sharded.solve_sharded_emulated([eng], rows)[0]

    once()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        r2c, _, st = once()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    if world > 1:
        tt = torch.tensor([ms], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    r2c = r2c.cpu().numpy()
    ok = bool((r2c >= 0).all() and len(set(r2c.tolist())) == rows)
    return {"workload": f"c5: dense {rows}x{cols} fp32 column-sharded over {world} rank(s), seed {SEED}, device-generated",
            "ms_per_solve": ms, "rounds": st["rounds"], "bids": st["bids"],
            "bid_gbs_all_ranks": st["bids"] * cols * 4 / ms / 1e6,
            "exchange": "all-gather of 32-byte records per round (NCCL)" if world > 1 else "none (single shard)",
            "exchanged_bytes_per_solve": st.get("exchanged_bytes", 0), "one_to_one_assignment": ok}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sharded", action="store_true", help="skip the column-sharded C5 leg")
    ap.add_argument("--no-sparse", action="store_true", help="skip the sparse (CSR) side measurement")
    ap.add_argument("--mht-problems", type=int, default=1 << 20, help="C3 problems per GPU")
    ap.add_argument("--c5-rows", type=int, default=C5["rows"])
    ap.add_argument("--c5-cols", type=int, default=C5["cols"])
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from auctionalgorithm_b200 import build
    build.build()
    import auctionalgorithm_b200 as ab

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        # rank 0 prints ONE JSON line on stdout: keep NCCL's version banner off it
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    w = WORKLOADS[args.workload]
    rows, cols = w["rows"], w["cols"]
    handle = ab.Handle(local)
    handle.set_stream(torch.cuda.current_stream().cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        tt = torch.tensor([x], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    # ---- device-resident solve --------------------------------------------------------------
    a_dev = torch.empty((rows, cols), dtype=torch.float32, device="cuda")
    r2c_dev = torch.empty(rows, dtype=torch.int32, device="cuda")
    handle.generate_device(a_dev.data_ptr(), 1, rows, cols, cols, ab.ROW_MAJOR, SEED)
    for _ in range(args.warmup):
        handle.solve_device(a_dev.data_ptr(), rows, cols, cols, ab.ROW_MAJOR, r2c_dev.data_ptr())
    sampler = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if rank == 0:
        sampler.start()
    e0.record()
    launches = 0
    for _ in range(args.steps):
        st = handle.solve_device(a_dev.data_ptr(), rows, cols, cols, ab.ROW_MAJOR, r2c_dev.data_ptr())
        launches += st["launches"]
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_step = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    solve_stats = st

    # ---- parity of what was just timed (rank 0, against the reference-made golden vector) ------
    parity_note = None
    gpath = os.path.join(ROOT, "tests", "golden", f"auction_{rows}x{cols}_s{SEED}.npz")
    if rank == 0 and os.path.exists(gpath):
        g = np.load(gpath)
        ref = np.full(rows, -1, np.int64)
        ref[g["x"]] = g["y"]
        parity_note = "exact" if np.array_equal(r2c_dev.cpu().numpy(), ref) else "differs-from-golden"

    # ---- bidding kernel alone: the roofline ---------------------------------------------------
    # Two states: the one the solves above left on the handle (prices and the rows' candidate lists: what the
    # first sweep of every eps phase after the first starts from; 7 of the 8 full sweeps of this solve) and
    # the one a solve starts from (zero prices, no candidates: the very first sweep).  The roofline entry is
    # the former; the latter is reported beside it.
    sweep_ms, first_ms = [], []
    for _ in range(3 + 5):
        ms, nbytes = handle.bid_sweep_device(a_dev.data_ptr(), rows, cols, cols, use_last_prices=True)
        sweep_ms.append(ms)
    for _ in range(3 + 5):
        ms, nbytes = handle.bid_sweep_device(a_dev.data_ptr(), rows, cols, cols, use_last_prices=False)
        first_ms.append(ms)
    sweep = statistics.mean(sweep_ms[3:])
    first_sweep = statistics.mean(first_ms[3:])
    peak, peak_src = peaks()
    # north_star's third quantity is the HBM bandwidth of the BIDDING kernel, the one kernel of the solve that
    # streams the matrix (one fp32 row per bid); the rest of a square solve is latency-bound narrow rounds
    # (k_narrow, one SM, reported as us_per_round / narrow_rounds in config), which no bandwidth roofline bounds
    roofline = {"bound": "hbm", "kernel": "k_bid<float> (bidding phase of a wide round, every row bids once; state left by "
                                          "the timed solves: their prices and candidate lists)",
                "achieved": nbytes / sweep / 1e6, "peak": peak, "unit": "GB/s",
                "frac": nbytes / sweep / 1e6 / peak, "traffic": ncu_traffic(), "peak_source": peak_src,
                "ms_per_launch": sweep, "algorithmic_bytes_per_launch": nbytes,
                "first_sweep": {"what": "same kernel from a solve's starting state (zero prices, no candidate lists)",
                                "ms_per_launch": first_sweep, "achieved": nbytes / first_sweep / 1e6,
                                "frac": nbytes / first_sweep / 1e6 / peak}}
    solve_gbs = solve_stats["bid_bytes"] / (ms_step * 1e6)

    # ---- end to end through the host API --------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        host = torch.empty((cols, rows), dtype=torch.float32).pin_memory()   # column-major storage (Eigen)
        host.copy_(a_dev.t())
        a_host = host.numpy().T                                             # F-contiguous rows x cols view
        del a_dev
        for _ in range(max(1, min(args.warmup, 2))):
            handle.solve(a_host)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            handle.solve(a_host)
        torch.cuda.synchronize()
        e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / args.steps)
        e2e = {"value": e2e_ms / world, "unit": "ms", "h2d_bytes_per_step": rows * cols * 4,
               "d2h_bytes_per_step": rows * 8, "ms_upload": handle.last_stats["ms_upload"],
               "ms_solve": handle.last_stats["ms_solve"], "ms_download": handle.last_stats["ms_download"],
               "input": "pinned host, column-major (Eigen default)"}
        del host, a_host
    else:
        del a_dev
    torch.cuda.empty_cache()

    # ---- C3: batched 32x48 problems, sharded over ranks, no collective -------------------------
    mht_ms, mht_bytes, mht_extra = run_mht(handle, torch, args.mht_problems, rank, max(args.steps, 3), 3,
                                           with_host_legs=(rank == 0 and world == 1 and not args.no_cpu))
    mht_ms = max_over_ranks(mht_ms)
    mht = {"problems_per_s": args.mht_problems * world / (mht_ms * 1e-3), "problems": args.mht_problems * world,
           "ms_per_step": mht_ms, "gbs_per_gpu": mht_bytes / mht_ms / 1e6,
           "frac_of_hbm_peak": mht_bytes / mht_ms / 1e6 / peak,
           "layout": "row-major 32x48 fp32 problems, device-generated, resident in HBM; one shard per rank",
           "kernel": "k_auction_batch_filter<6,true,4,3>", **mht_extra}
    torch.cuda.empty_cache()

    # ---- C5: column-sharded big problem ------------------------------------------------------------
    shard = None
    if not args.no_sharded:
        shard = run_sharded(handle, torch, dist, rank, world, 2, args.c5_rows, args.c5_cols)

    # ---- sparse rows: the reference's SparseMatrix<Scalar, RowMajor> instantiation --------
    sparse = None
    if rank == 0 and not args.no_sparse:
        sparse = run_sparse(handle)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        arm = CpuArm(rows, cols)
        cpu = arm.sample(12.0)
        cpu["auctionmt"] = arm.mt_note()

    if rank == 0:
        big = rows * cols * 4 > 126e6
        print(json.dumps({
            "metric": f"solve_ms_{rows}x{cols}_fp32", "value": ms_step / world, "unit": "ms", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": False,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": (f"{args.workload}: dense {rows}x{cols} fp32 (BASELINE.json configs), uniform grid "
                                    f"(0,1], seed {SEED}, device-generated; "
                                    + ("inputs larger than L2" if big else "L2-resident")),
                       "parallelism": "single GPU" if world == 1 else f"{world} independent replicas (a mid-size dense solve does not shard)",
                       "eps_final": solve_stats["eps_final"], "rounds": solve_stats["rounds"], "bids": solve_stats["bids"],
                       "phases": solve_stats["phases"], "narrow_rounds": solve_stats["narrow_rounds"],
                       "one_bidder_rounds": solve_stats["tail_bids"], "schedule_restarts": solve_stats["tail_runs"],
                       "cache_hits": solve_stats["cache_hits"], "cache_misses": solve_stats["cache_misses"],
                       "eps_first": solve_stats["eps_first"],
                       "us_per_round": 1e3 * solve_stats["ms_solve"] / max(solve_stats["rounds"], 1),
                       "parity_vs_reference_golden": parity_note},
            "roofline": roofline,
            "solve_algorithmic_gbs": solve_gbs,
            "cpu_baseline": cpu,
            "e2e": e2e,
            "mht": mht,
            "sharded": shard,
            "sparse": sparse,
            "gpu_launches": launches,
            "clocks": clocks,
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
