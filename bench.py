#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 auction path (contract: task brief §4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

BASELINE.json's metric has three quantities: solve ms at 16384 x 16384 fp32, batched 32x48 MHT problems/s, and the
bidding kernel's HBM GB/s against the peak.  All three are in every line; which one is `value` depends on N:

  N = 1   `value` = C4, the configuration the metric is quoted on: one dense 16384 x 16384 fp32 problem (1 GiB, larger
          than L2) solved to the reference's assignment, ms per solve with the matrix resident in HBM.  A step = one
          full solve.  `e2e` = the same solve through the host API (pinned column-major host matrix in, row->column
          out; H2D + transpose + solve + D2H inside the timed call).  `roofline` = the bidding kernel from a solve's
          starting state (the cold sweep every solve pays), the warm one beside it, and `roofline_step` for the whole
          solve.  Also in the line: `c1`, `c2` (BASELINE configs 0 and 1, GPU and fully measured CPU), `mht` (C3 on the
          eight shards the ranks of an 8-GPU run get), `sharded` (C5 on one GPU), `sparse`, `cpu_baseline`.
  N > 1   a single mid-size dense problem does not shard ("replicas only"), so `value` = the path that
          does: C3, 2^20 independent 32x48 problems per GPU, problems/s over all ranks (`scaling: weak`, no
          collective on the data path; max over ranks of the CUDA-event time).  A step = one pass over every rank's
          shard.  `mht.per_rank` lists every rank's ms per step, takeovers and largest bid count.  `c4_replicas` =
          every rank solving its own copy of C4 (ms per solve, NOT divided by N), `sharded` = C5 column-sharded over
          the ranks with one exchange per round.

`--impl reference` times the reference's own CPU implementation on the box's host cores for the same `value`:
N = 1: the unmodified reference (oracle/_ref) solved ONCE on the largest square that fits the arm's time
(`measured_full`), and C4 itself as an extrapolation from a bounded sample of bids, flagged `extrapolated` with the
method's error where both exist; N > 1: the unmodified reference's Auction<double>::solve looped over a sample of
the same MHT problems on all host threads.
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

C1 = dict(rows=200, cols=300)
C2 = dict(rows=2048, cols=3072)
C4 = dict(rows=16384, cols=16384)
C5 = dict(rows=65536, cols=98304)
MHT_ROWS, MHT_COLS = 32, 48
MHT_BYTES = MHT_ROWS * MHT_COLS * 4 + MHT_ROWS * 4     # algorithmic bytes per problem: matrix in, int32 row->col out
SEED = 1
READ_ONLY_PEAK_GBS = 7000.0   # what a read-only stream reaches on this part (tools/micro/read_probe.cu, round 1)


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


def profile_number(name: str, key: str):
    """A per-launch figure from a committed ncu --set full summary under profiles/."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", name)))[key]
    except Exception:
        return None


def c4_reference_counters() -> dict:
    p = os.path.join(ROOT, "tests", "golden", "c4_oracle_stats.json")
    return json.load(open(p)) if os.path.exists(p) else {}


def host_cpu() -> dict:
    model = None
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    return {"model": model, "logical_cpus": os.cpu_count() or 1}


# ======================================================================================================
# CPU side: the reference on the host cores (oracle/_ref = the unmodified reference headers compiled by
# oracle/Makefile; oracle/ = the C port, bit-identical to it, which also accepts a bid budget)
# ======================================================================================================
class CpuArm:
    def __init__(self):
        import numpy as np
        from oracle import pyoracle as po
        po.build()
        self.po, self.np = po, np
        self.cores = os.cpu_count() or 1

    def matrix(self, rows, cols):
        return self.np.asfortranarray(self.po.generate(rows, cols, SEED).astype(self.np.float64))  # Eigen layout

    # ---- small configs: everything measured in full ------------------------------------------------
    def small_config(self, rows, cols, mt_threads) -> dict:
        """BASELINE.md §3 table for a config the CPU finishes in seconds: Auction<double>, Auction<float, MatrixXf>
        and AuctionMT<double> (flagged: it solves the leading rows x rows block when rows < cols)."""
        po, np = self.po, self.np
        a32 = po.generate(rows, cols, SEED)
        a64 = np.asfortranarray(a32.astype(np.float64))
        out = {"kind": "reference" if po.reference() is not None else "port", "host": host_cpu()}
        if po.reference() is None:
            t0 = time.perf_counter()
            po.oracle_solve(a64)
            out["auction_double_ms"] = (time.perf_counter() - t0) * 1e3
            return out
        out["auction_double_ms"] = min(po.reference_solve(a64, "auction")["seconds"] for _ in range(3)) * 1e3
        if rows != cols and rows * cols <= 2048 * 3072:   # Auction<float> livelocks on square / large inputs
            out["auction_float_ms"] = min(po.reference_solve(np.asfortranarray(a32), "auction")["seconds"]
                                          for _ in range(3)) * 1e3
        mt = {}
        for t in mt_threads:
            mt[str(t)] = po.reference_solve(a64, "mt", n_threads=t)["seconds"] * 1e3
        out["auctionmt_double_ms_by_threads"] = mt
        if rows < cols:
            out["auctionmt_note"] = "AuctionMT solves only the leading rows x rows block when rows < cols (SURVEY §0.2): timing only"
        return out

    # ---- squares: one full measured solve, AuctionMT sweep ----------------------------------------------
    def measured_full(self, n: int) -> dict:
        po = self.po
        a = self.matrix(n, n)
        if po.reference() is not None:
            sec = po.reference_solve(a, "auction")["seconds"]
            return {"n": n, "seconds": sec, "kind": "reference", "what": f"Auction<double>::solve on {n}x{n} seed {SEED}, "
                    "one run of the unmodified reference (Auction.h:44), single thread"}
        t0 = time.perf_counter()
        po.oracle_solve(a)
        return {"n": n, "seconds": time.perf_counter() - t0, "kind": "port", "what": f"CPU port on {n}x{n} seed {SEED}"}

    def auctionmt_sweep(self, n: int) -> dict:
        """AuctionMT<double> (AuctionMT.h:50) at 1, 2, 4, ... threads up to the host's logical CPUs."""
        po = self.po
        if po.reference() is None:
            return {"unavailable": "oracle/_ref not built"}
        a = self.matrix(n, n)
        st = po.reference_solve(a, "auction")["seconds"]
        ts, t = [], 1
        while t < self.cores:
            ts.append(t)
            t *= 2
        ts.append(self.cores)
        by = {}
        for t in ts:
            by[str(t)] = po.reference_solve(a, "mt", n_threads=t)["seconds"] * 1e3
        best = min(by, key=lambda k: by[k])
        return {"case": f"{n}x{n} double, seed {SEED}", "auction_single_thread_ms": st * 1e3, "auctionmt_ms_by_threads": by,
                "best_threads": int(best), "best_ms": by[best], "all_cores_threads": self.cores,
                "all_cores_ms": by[str(self.cores)], "host": host_cpu()}

    # ---- C4: too long to run (21 min): bounded sample of bids, scaled, flagged ---------------------------
    def extrapolate(self, rows, cols, total_bids, seconds_target) -> dict:
        """First bids of Auction<double> (CPU port with a bid budget, bit-identical to the reference) scaled to
        total_bids.  Per-bid cost is not constant over a solve (the edge scan of Auction.h:264-285 grows), so this is
        an estimate; `extrapolation_error_at_measured_n` says how far the same method is off where the truth is known."""
        po = self.po
        a = self.matrix(rows, cols)
        po.oracle_solve(a, max_bids=500)
        t0 = time.perf_counter()
        probe = po.oracle_solve(a, max_bids=2000)
        per_bid = (time.perf_counter() - t0) / max(probe["stats"]["bids"], 1)
        budget = max(2000, int(seconds_target / per_bid))
        t0 = time.perf_counter()
        out = po.oracle_solve(a, max_bids=budget)
        sec = time.perf_counter() - t0
        bids = out["stats"]["bids"]
        done = out["status"] == po.ORC_OK
        return {"sample_bids": bids, "sample_seconds": sec, "total_bids": bids if done else total_bids,
                "seconds": sec if done else sec / bids * total_bids, "complete": done}

    # ---- MHT: the reference has no batched entry, solve() in a loop ------------------------------------------
    def mht(self, host_batch, gpu_r2c=None) -> dict:
        po, np = self.po, self.np
        n = len(host_batch)
        out = {"cores": self.cores, "sample": f"first {n} problems of rank 0's shard", "host": host_cpu()}
        if po.reference() is not None and hasattr(po.reference(), "ref_auction_batch_f32_as_f64"):
            r1, s1 = po.reference_batch(host_batch[:n // 8], 1)
            rN, sN = po.reference_batch(host_batch, self.cores)
            out.update(kind="reference", double_1thread_problems_per_s=(n // 8) / s1,
                       double_all_threads_problems_per_s=n / sN,
                       what="the unmodified reference's Auction<double>::solve (Auction.h:44) once per problem")
            ref = rN
        else:
            t0 = time.perf_counter()
            ref, failed = po.oracle_batch(host_batch, use_double=True, n_threads=self.cores)
            out.update(kind="port", double_all_threads_problems_per_s=n / (time.perf_counter() - t0))
        t0 = time.perf_counter()
        po.oracle_batch(host_batch, use_double=False, n_threads=self.cores)
        out["float_port_all_threads_problems_per_s"] = n / (time.perf_counter() - t0)
        if gpu_r2c is not None:
            out["gpu_result_identical"] = bool((ref == gpu_r2c[:n]).all())
        return out


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    arm = CpuArm()
    reps = args.warmup + args.steps
    if args.gpus > 1:
        # the path that shards: MHT problems/s on all host threads, a bounded sample per step
        np = arm.np
        n = 1 << 15
        host = np.stack([arm.po.generate(MHT_ROWS, MHT_COLS, SEED, problem=b) for b in range(n)])
        vals, last = [], None
        for k in range(reps):
            last = arm.mht(host if k == reps - 1 else host[:n // 4])
            if k >= args.warmup:
                vals.append(last["double_all_threads_problems_per_s"])
        v = statistics.mean(vals)
        print(json.dumps({
            "impl": "reference", "metric": "mht_32x48_problems_per_s", "value": v, "unit": "problems/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * (n // 4) / v,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"c3: batched MHT data association, 32x48 fp32 problems, seed {SEED}; the reference's "
                                   "Auction<double>::solve once per problem on all host threads (it has no batched "
                                   "entry and no GPU path: the host does not grow with --gpus); a step = a bounded sample of "
                                   f"{n // 4} problems"},
            "cpu_baseline": dict(last, value=v, unit="problems/s"),
            "e2e": {"value": v, "unit": "problems/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }))
        return
    ref = c4_reference_counters()
    full_n = args.ref_full_n
    full = arm.measured_full(full_n)
    # the same estimate where the truth was just measured: its error is a number, not a hope
    g = os.path.join(ROOT, "tests", "golden", f"auction_{full_n}x{full_n}_s{SEED}.npz")
    err = None
    if os.path.exists(g):
        port_full = arm.extrapolate(full_n, full_n, 0, 1e9)                        # the port, whole solve
        est = arm.extrapolate(full_n, full_n, port_full["sample_bids"], max(2.0, 0.01 * port_full["sample_seconds"]))
        err = {"n": full_n, "predicted_s": est["seconds"], "port_full_s": port_full["sample_seconds"],
               "reference_full_s": full["seconds"], "predicted_over_reference": est["seconds"] / full["seconds"],
               "sample_bids": est["sample_bids"], "total_bids": port_full["sample_bids"]}
    vals, last = [], None
    per = min(6.0, max(1.0, 60.0 / reps))
    for k in range(reps):
        last = arm.extrapolate(C4["rows"], C4["cols"], ref.get("bids", 0), per)
        if k >= args.warmup:
            vals.append(last["seconds"] * 1e3)
    v = statistics.mean(vals)
    cpu = {"value": v, "unit": "ms", "cores": 1, "kind": "port", "extrapolated": True,
           "sample": (f"first {last['sample_bids']} of the {last['total_bids']} bids of Auction<double>::solve on C4 "
                      f"(CPU port with a bid budget, bit-identical to the reference) in {last['sample_seconds']:.1f} s, "
                      "scaled by the bid count; the unmodified reference measured offline took "
                      f"{ref.get('seconds_reference', 0):.0f} s (tests/golden/c4_oracle_stats.json)"),
           "measured_full": full, "extrapolation_error_at_measured_n": err, "host": host_cpu(),
           "note": "Auction::solve is single-threaded by construction (Auction.h:44-109); AuctionMT is slower at every "
                   "size measured (cpu_baseline.auctionmt_sweep in the b200 arm's line)"}
    print(json.dumps({
        "impl": "reference", "metric": f"solve_ms_{C4['rows']}x{C4['cols']}_fp32", "value": v, "unit": "ms",
        "extrapolated": True, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": v,
        "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"c4: dense {C4['rows']}x{C4['cols']} fp32 widened to f64, seed {SEED}: the reference's "
                               "Auction<double>::solve on the host, single thread.  `value` is an EXTRAPOLATION from a "
                               "bounded sample of bids (a full solve takes ~21 min); `cpu_baseline.measured_full` is a "
                               "real, complete run of the unmodified reference on the largest square that fits this arm"},
        "cpu_baseline": cpu,
        "e2e": {"value": v, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ======================================================================================================
# GPU side
# ======================================================================================================
def time_dense(handle, torch, ab, rows, cols, steps, warmup, barrier=None):
    """Device-resident solve: (ms per solve by CUDA events over `steps` solves, last stats, r2c tensor, matrix)."""
    a_dev = torch.empty((rows, cols), dtype=torch.float32, device="cuda")
    r2c_dev = torch.empty(rows, dtype=torch.int32, device="cuda")
    handle.generate_device(a_dev.data_ptr(), 1, rows, cols, cols, ab.ROW_MAJOR, SEED)
    for _ in range(warmup):
        handle.solve_device(a_dev.data_ptr(), rows, cols, cols, ab.ROW_MAJOR, r2c_dev.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if barrier:
        barrier()
    torch.cuda.synchronize()
    e0.record()
    launches = 0
    for _ in range(steps):
        st = handle.solve_device(a_dev.data_ptr(), rows, cols, cols, ab.ROW_MAJOR, r2c_dev.data_ptr())
        launches += st["launches"]
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, st, r2c_dev, a_dev, launches


def golden_parity(np, rows, cols, r2c_dev):
    gpath = os.path.join(ROOT, "tests", "golden", f"auction_{rows}x{cols}_s{SEED}.npz")
    if not os.path.exists(gpath):
        return None
    g = np.load(gpath)
    ref = np.full(rows, -1, np.int64)
    ref[g["x"]] = g["y"]
    return "exact" if np.array_equal(r2c_dev.cpu().numpy(), ref) else "differs-from-golden"


def time_e2e_dense(handle, torch, a_dev, rows, cols, steps, warmup):
    """The host API: pinned column-major (Eigen default) matrix in, row->column out; copies inside the timed call."""
    host = torch.empty((cols, rows), dtype=torch.float32).pin_memory()
    host.copy_(a_dev.t())
    a_host = host.numpy().T                                             # F-contiguous rows x cols view
    for _ in range(max(1, min(warmup, 2))):
        handle.solve(a_host)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        handle.solve(a_host)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    st = handle.last_stats
    return {"value": ms, "unit": "ms", "h2d_bytes_per_step": rows * cols * 4, "d2h_bytes_per_step": rows * 8,
            "ms_upload": st["ms_upload"], "ms_solve": st["ms_solve"], "ms_download": st["ms_download"],
            "input": "pinned host, column-major (Eigen default)"}


def small_config(handle, torch, ab, np, name, rows, cols, cpu_arm):
    """c1 / c2: BASELINE configs 0 and 1.  Latency per solve and per auction round (the small-problem regime)."""
    ms, st, r2c, a_dev, _ = time_dense(handle, torch, ab, rows, cols, 20, 5)
    e2e = time_e2e_dense(handle, torch, a_dev, rows, cols, 20, 3)
    out = {"workload": f"{name}: dense {rows}x{cols} fp32, seed {SEED}" + ("" if rows * cols * 4 > 126e6 else " (L2-resident)"),
           "gpu_ms_device_resident": ms, "gpu_ms_kernels_only": st["ms_solve"], "gpu_ms_e2e_host_matrix": e2e["value"],
           "rounds": st["rounds"], "bids": st["bids"], "phases": st["phases"],
           "us_per_round": 1e3 * st["ms_solve"] / max(st["rounds"], 1), "launches_per_solve": st["launches"],
           "parity_vs_reference_golden": golden_parity(np, rows, cols, r2c)}
    if cpu_arm is not None:
        threads = sorted({1, 2, 4, cpu_arm.cores}) if rows * cols <= 2048 * 3072 else [2]
        out["cpu"] = cpu_arm.small_config(rows, cols, threads)
    del a_dev
    return out


def run_mht_shard(handle, torch, n_problems: int, problem0: int, steps: int, warmup: int, t=None, out=None):
    """One shard of C3, generated on the device (inputs resident in HBM).  Returns (ms per step by CUDA events over
    `steps` launches, mean kernel ms, last stats, launches)."""
    if t is None:
        t = torch.empty((n_problems, MHT_ROWS, MHT_COLS), dtype=torch.float32, device="cuda")
        out = torch.empty((n_problems, MHT_ROWS), dtype=torch.int32, device="cuda")
    handle.generate_device(t.data_ptr(), n_problems, MHT_ROWS, MHT_COLS, MHT_COLS, 1, SEED, problem0=problem0)
    for _ in range(warmup):
        handle.solve_batch_device(t.data_ptr(), n_problems, MHT_ROWS, MHT_COLS, out.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    kernel_ms, launches = 0.0, 0
    for _ in range(steps):
        st = handle.solve_batch_device(t.data_ptr(), n_problems, MHT_ROWS, MHT_COLS, out.data_ptr())
        kernel_ms += st["ms_solve"]
        launches += st["launches"]
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, kernel_ms / steps, st, launches, t, out


def mht_e2e(handle, torch, t, n_e2e: int):
    """C3 through the reference-facing batched entry (b200lsap_solve_batch_f32): HOST arrays in and out, the library's
    own copies inside the timed call."""
    host_t = torch.empty((n_e2e, MHT_ROWS, MHT_COLS), dtype=torch.float32, pin_memory=True)
    host_t.copy_(t[:n_e2e])
    torch.cuda.synchronize()
    host = host_t.numpy()
    handle.solve_batch(host[:1024])
    handle.solve_batch(host)
    t0 = time.perf_counter()
    r2c = handle.solve_batch(host)
    sec = time.perf_counter() - t0
    assert (r2c >= 0).all()
    return host, r2c, {"problems_per_s": n_e2e / sec, "problems": n_e2e, "h2d_bytes": int(host.nbytes),
                       "d2h_bytes": int(r2c.nbytes), "ms": sec * 1e3, "h2d_gbs": host.nbytes / sec / 1e9,
                       "input": "pinned host array [batch, 32, 48] fp32, pageable result array; bound by the 6144 "
                                "bytes per problem that cross PCIe, not by the kernel"}


def mht_frames(handle, torch, t, out_t):
    """Latency of small MHT frames (north_star: the small-problem regime): device-resident batches of 64 and 1024."""
    res = {}
    for n in (64, 1024):
        for _ in range(5):
            handle.solve_batch_device(t.data_ptr(), n, MHT_ROWS, MHT_COLS, out_t.data_ptr())
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = 50
        k = 0.0
        for _ in range(reps):
            k += handle.solve_batch_device(t.data_ptr(), n, MHT_ROWS, MHT_COLS, out_t.data_ptr())["ms_solve"]
        res[str(n)] = {"us_per_call_wall": (time.perf_counter() - t0) / reps * 1e6, "us_kernel": k / reps * 1e3}
    return res


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


def run_sharded(handle, torch, dist, rank: int, world: int, steps: int, rows: int, cols: int):
    """C5: one rows x cols problem column-sharded over the ranks, generated per shard on the device."""
    from auctionalgorithm_b200 import sharded
    return sharded.bench_leg(handle, torch, dist, rank, world, steps, rows, cols, SEED)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip every CPU leg (cpu_baseline, c1/c2 CPU times)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sharded", action="store_true", help="skip the column-sharded C5 leg")
    ap.add_argument("--no-sparse", action="store_true", help="skip the sparse (CSR) side measurement")
    ap.add_argument("--no-small", action="store_true", help="skip c1 / c2")
    ap.add_argument("--mht-problems", type=int, default=1 << 20, help="C3 problems per GPU")
    ap.add_argument("--mht-shards", type=int, default=8, help="N = 1: how many of the 8-GPU run's shards to time")
    ap.add_argument("--c5-rows", type=int, default=C5["rows"])
    ap.add_argument("--c5-cols", type=int, default=C5["cols"])
    ap.add_argument("--ref-full-n", type=int, default=4096, help="--impl reference: square solved once in full")
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
        # rank 0 prints ONE JSON line on stdout: whatever NCCL logs (version banner included) goes to a file
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/nccl.%h.%p.log")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    handle = ab.Handle(local)
    handle.set_stream(torch.cuda.current_stream().cuda_stream)
    peak, peak_src = peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        tt = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    def gather(vals):
        tt = torch.tensor(vals, device="cuda", dtype=torch.float64)
        if world == 1:
            return [tt.tolist()]
        outl = [torch.empty_like(tt) for _ in range(world)]
        dist.all_gather(outl, tt)
        return [o.tolist() for o in outl]

    if world == 1:
        return main_single(args, handle, torch, dist, ab, np, peak, peak_src)

    # ================================ N > 1: the path that shards ===========================================
    n = args.mht_problems
    steps = max(args.steps, 3)
    # warm-up outside the clock (also allocates), then the timed K steps between barriers
    ms0, _, _, _, t, out = run_mht_shard(handle, torch, n, rank * n, 1, max(args.warmup, 3))
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    ms, kernel_ms, st, launches, t, out = run_mht_shard(handle, torch, n, rank * n, steps, 0, t, out)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    per_rank = gather([ms, kernel_ms, st["batch_takeovers"], st["batch_max_bids"]])
    ms_step = max(p[0] for p in per_rank)
    value = n * world / (ms_step * 1e-3)
    # end to end: every rank feeds its GPU from its own pinned host batch at the same time
    e2e = None
    if not args.no_e2e:
        n_e2e = min(n, 1 << 18)
        barrier()
        _, r2c_h, e = mht_e2e(handle, torch, t, n_e2e)
        sec = max_over_ranks(e["ms"] * 1e-3)
        e2e = {"value": n_e2e * world / sec, "unit": "problems/s", "h2d_bytes_per_step": e["h2d_bytes"],
               "d2h_bytes_per_step": e["d2h_bytes"], "problems_per_rank": n_e2e, "ms": sec * 1e3, "input": e["input"]}
    del t, out
    torch.cuda.empty_cache()
    # replicas of C4: what a dense mid-size solve does on N GPUs (nothing: every rank solves its own copy)
    c4_ms, c4_st, c4_r2c, a_dev, _ = time_dense(handle, torch, ab, C4["rows"], C4["cols"], 3, 2, barrier)
    c4_par = golden_parity(np, C4["rows"], C4["cols"], c4_r2c)
    c4_all = gather([c4_ms])
    del a_dev, c4_r2c
    torch.cuda.empty_cache()
    shard = None
    if not args.no_sharded:
        shard = run_sharded(handle, torch, dist, rank, world, 2, args.c5_rows, args.c5_cols)
    if rank == 0:
        kernel_step = max(p[1] for p in per_rank)
        print(json.dumps({
            "metric": "mht_32x48_problems_per_s", "value": value, "unit": "problems/s", "n_gpus": world,
            "steps": steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"c3: batched MHT data association (BASELINE.json configs[2]), {n} independent 32x48 fp32 "
                                   f"problems per GPU, seed {SEED}, rank r holds problems [r*{n}, (r+1)*{n}), device-generated, "
                                   "resident in HBM; 6.4 GB per GPU per step (larger than L2)",
                       "parallelism": f"{world} ranks, one shard each, no data-path collective",
                       "input_dtype": "f32", "arithmetic": "fp32 filter, exact fp64 bids (Auction<double> on the widened data)"},
            "roofline": {"bound": "hbm", "kernel": "k_auction_batch_filter<6,true,4,4> (the step is this one kernel)",
                         "achieved": n * MHT_BYTES / kernel_step / 1e6, "peak": peak, "unit": "GB/s",
                         "frac": n * MHT_BYTES / kernel_step / 1e6 / peak, "peak_source": peak_src,
                         "traffic": profile_number("r2_ncu_batch_summary.json", "dram_bytes_per_launch"),
                         "algorithmic_bytes_per_launch": n * MHT_BYTES, "ms_per_launch": kernel_step,
                         "note": "slowest rank's mean launch, CUDA events on the launching stream"},
            "cpu_baseline": None,
            "e2e": e2e,
            "mht": {"problems_per_s": value, "problems": n * world, "ms_per_step": ms_step,
                    "per_rank": [{"rank": r, "ms_per_step": p[0], "kernel_ms": p[1], "takeovers": int(p[2]),
                                  "max_bids_per_problem": int(p[3])} for r, p in enumerate(per_rank)],
                    "frac_of_hbm_peak_slowest_rank": n * MHT_BYTES / kernel_step / 1e6 / peak},
            "c4_replicas": {"what": "every rank solves its own copy of C4 (a mid-size dense solve does not shard); "
                                    "ms per solve, not divided by the number of ranks",
                            "ms_per_solve_by_rank": [c[0] for c in c4_all], "rounds": c4_st["rounds"],
                            "parity_vs_reference_golden": c4_par},
            "sharded": shard,
            "gpu_launches": launches,
            "clocks": clocks,
        }))
    dist.destroy_process_group()


def main_single(args, handle, torch, dist, ab, np, peak, peak_src):
    rows, cols = C4["rows"], C4["cols"]
    cpu_arm = None if args.no_cpu else CpuArm()
    # ---- device-resident solve (the headline) ------------------------------------------------------------
    sampler = ClockSampler(torch.cuda.current_device())
    # (time_dense warms up, synchronises, then times `steps` solves with CUDA events)
    a_tmp = None
    handle_ms = None
    sampler.start()
    ms_step, solve_stats, r2c_dev, a_dev, launches = time_dense(handle, torch, ab, rows, cols, args.steps, args.warmup)
    clocks = sampler.stop()
    parity_note = golden_parity(np, rows, cols, r2c_dev)

    # ---- bidding kernel alone: the roofline ------------------------------------------------------------------
    # Two states: the one a solve starts from (zero prices, no candidate lists: the cold sweep every solve pays) and
    # the one the solves above left on the handle (their prices and the rows' candidate lists: what a full sweep of a
    # later eps phase starts from when it has to scan, i.e. when the row caches were not reusable).
    warm_ms, cold_ms = [], []
    for _ in range(3 + 5):
        ms, nbytes = handle.bid_sweep_device(a_dev.data_ptr(), rows, cols, cols, use_last_prices=True)
        warm_ms.append(ms)
    for _ in range(3 + 5):
        ms, nbytes = handle.bid_sweep_device(a_dev.data_ptr(), rows, cols, cols, use_last_prices=False)
        cold_ms.append(ms)
    warm, cold = statistics.mean(warm_ms[3:]), statistics.mean(cold_ms[3:])
    roofline = {"bound": "hbm", "kernel": "k_bid<float>: the bidding phase of a wide round, every row bids once, from a "
                                          "solve's starting state (zero prices, no candidate lists)",
                "achieved": nbytes / cold / 1e6, "peak": peak, "unit": "GB/s", "frac": nbytes / cold / 1e6 / peak,
                "traffic": profile_number("kbid_ncu_summary.json", "dram_bytes_per_launch"), "peak_source": peak_src,
                "ms_per_launch": cold, "algorithmic_bytes_per_launch": nbytes,
                "frac_of_read_only_peak": nbytes / cold / 1e6 / READ_ONLY_PEAK_GBS,
                "read_only_peak_gbs": READ_ONLY_PEAK_GBS,
                "warm_sweep": {"what": "same kernel on the prices and candidate lists a solve leaves (kept for comparison: since the row caches are reused across phases the later full sweeps of a solve no longer scan the matrix at all)",
                               "ms_per_launch": warm, "achieved": nbytes / warm / 1e6, "frac": nbytes / warm / 1e6 / peak,
                               "frac_of_read_only_peak": nbytes / warm / 1e6 / READ_ONLY_PEAK_GBS},
                "note": "the copy peak counts read + write bytes; a read-only stream reaches ~7.0 TB/s on this part, so "
                        "frac can exceed 1 against the copy peak; these sweeps are ~4 % of the timed step, see roofline_step"}
    roofline_step = {"what": "whole solve: algorithmic bytes of all bids (bids x cols x 4) / ms_per_step / peak; the solve is "
                             "round-latency-bound (narrow rounds on one SM), not bandwidth-bound",
                     "bid_bytes": solve_stats["bid_bytes"], "gbs": solve_stats["bid_bytes"] / (ms_step * 1e6),
                     "frac": solve_stats["bid_bytes"] / (ms_step * 1e6) / peak}

    # ---- end to end through the host API --------------------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        e2e = time_e2e_dense(handle, torch, a_dev, rows, cols, args.steps, args.warmup)
    del a_dev, r2c_dev
    torch.cuda.empty_cache()

    # ---- c1, c2: the small-problem regime -----------------------------------------------------------------------
    c1 = c2 = None
    if not args.no_small:
        c1 = small_config(handle, torch, ab, np, "c1", C1["rows"], C1["cols"], cpu_arm)
        c2 = small_config(handle, torch, ab, np, "c2", C2["rows"], C2["cols"], cpu_arm)

    # ---- C3: batched 32x48 problems; the shards the ranks of an 8-GPU run get, one after the other ------------------
    n = args.mht_problems
    steps = max(args.steps, 3)
    shards, t, out = [], None, None
    for k in range(max(1, args.mht_shards)):
        ms, kms, st, _, t, out = run_mht_shard(handle, torch, n, k * n, steps, 3, t, out)
        shards.append({"shard": k, "problem0": k * n, "ms_per_step": ms, "kernel_ms": kms,
                       "problems_per_s": n / (ms * 1e-3), "takeovers": st["batch_takeovers"],
                       "max_bids_per_problem": st["batch_max_bids"]})
    s0 = shards[0]
    worst = max(s["ms_per_step"] for s in shards)
    best = min(s["ms_per_step"] for s in shards)
    mht = {"problems_per_s": s0["problems_per_s"], "problems": n, "ms_per_step": s0["ms_per_step"],
           "kernel_ms": s0["kernel_ms"], "gbs_per_gpu": n * MHT_BYTES / s0["kernel_ms"] / 1e6,
           "frac_of_hbm_peak": n * MHT_BYTES / s0["kernel_ms"] / 1e6 / peak,
           "traffic": profile_number("r2_ncu_batch_summary.json", "dram_bytes_per_launch"),
           "layout": "row-major 32x48 fp32 problems, device-generated, resident in HBM",
           "kernel": "k_auction_batch_filter<6,true,4,4>",
           "shards": shards, "worst_shard_over_best": worst / best,
           "slowest_shard_problems_per_s": n / (worst * 1e-3),
           "note": "problems_per_s is shard 0 (what rank 0 of an N-GPU run holds); `shards` = the same launch on the shards "
                   "of ranks 0..7, whose slowest problem needs 387 to 2148 bids (mean 44)"}
    # reload shard 0 for the host legs
    handle.generate_device(t.data_ptr(), n, MHT_ROWS, MHT_COLS, MHT_COLS, 1, SEED, problem0=0)
    mht["frame_latency"] = mht_frames(handle, torch, t, out)
    if not args.no_e2e:
        host, r2c_h, mht["e2e"] = mht_e2e(handle, torch, t, min(n, 1 << 18))
        if cpu_arm is not None:
            mht["cpu_baseline"] = cpu_arm.mht(host[:1 << 15], r2c_h)
    del t, out
    torch.cuda.empty_cache()

    # ---- C5: column-sharded big problem, here on one GPU (the same code path with a single shard) -------------------
    shard = None
    if not args.no_sharded:
        shard = run_sharded(handle, torch, dist, 0, 1, 2, args.c5_rows, args.c5_cols)

    # ---- sparse rows: the reference's SparseMatrix<Scalar, RowMajor> instantiation --------
    sparse = None if args.no_sparse else run_sparse(handle)

    # ---- the CPU beside it: bounded, labelled --------------------------------------------------------------------
    cpu = None
    if cpu_arm is not None:
        ref = c4_reference_counters()
        est = cpu_arm.extrapolate(rows, cols, ref.get("bids", 0), 10.0)
        sweep = cpu_arm.auctionmt_sweep(1024)
        cpu = {"value": est["seconds"] * 1e3, "unit": "ms", "cores": 1, "kind": "port", "extrapolated": True,
               "sample": (f"first {est['sample_bids']} of the {est['total_bids']} bids of Auction<double>::solve on C4 (CPU "
                          f"port, bit-identical to the reference, which has no bid budget) in {est['sample_seconds']:.1f} s, "
                          "scaled by the bid count; the unmodified reference measured offline took "
                          f"{ref.get('seconds_reference', 0):.0f} s (tests/golden/c4_oracle_stats.json); a complete measured "
                          "run of the unmodified reference is in `bench.py --impl reference` (measured_full)"),
               "host": host_cpu(), "auctionmt_sweep": sweep}
        if sweep.get("auction_single_thread_ms"):
            # north_star phrases the C4 target against AuctionMT on all cores: scale the single-thread C4 figure by
            # AuctionMT's measured ratio to Auction at 1024^2 (labelled: an extrapolation of an extrapolation)
            r_all = sweep["all_cores_ms"] / sweep["auction_single_thread_ms"]
            r_best = sweep["best_ms"] / sweep["auction_single_thread_ms"]
            cpu["c4_auctionmt_all_cores_ms_extrapolated"] = est["seconds"] * 1e3 * r_all
            cpu["c4_auctionmt_best_threads_ms_extrapolated"] = est["seconds"] * 1e3 * r_best
            cpu["gpu_fraction_of_auctionmt_best"] = ms_step / (est["seconds"] * 1e3 * r_best)

    big = rows * cols * 4 > 126e6
    print(json.dumps({
        "metric": f"solve_ms_{rows}x{cols}_fp32", "value": ms_step, "unit": "ms", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": False,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": (f"c4: dense {rows}x{cols} fp32 (BASELINE.json configs[3]), uniform grid (0,1], seed {SEED}, "
                                "device-generated; " + ("inputs larger than L2" if big else "L2-resident")),
                   "parallelism": "single GPU", "input_dtype": "f32",
                   "arithmetic": "fp32 filter, exact fp64 prices and bids (Auction<double> on the widened data)",
                   "eps_final": solve_stats["eps_final"], "rounds": solve_stats["rounds"], "bids": solve_stats["bids"],
                   "phases": solve_stats["phases"], "narrow_rounds": solve_stats["narrow_rounds"],
                   "one_bidder_rounds": solve_stats["tail_bids"], "schedule_restarts": solve_stats["tail_runs"],
                   "cache_hits": solve_stats["cache_hits"], "cache_misses": solve_stats["cache_misses"],
                   "eps_first": solve_stats["eps_first"],
                   "us_per_round": 1e3 * solve_stats["ms_solve"] / max(solve_stats["rounds"], 1),
                   "parity_vs_reference_golden": parity_note},
        "roofline": roofline,
        "roofline_step": roofline_step,
        "cpu_baseline": cpu,
        "e2e": e2e,
        "c1": c1,
        "c2": c2,
        "mht": mht,
        "sharded": shard,
        "sparse": sparse,
        "gpu_launches": launches,
        "clocks": clocks,
    }))


if __name__ == "__main__":
    main()
