"""Generates the golden vectors under tests/golden/ by running the UNMODIFIED reference
(oracle/_ref/libref_auction.so, built by `make -C oracle` from /root/reference).

The reference ships no tests, fixtures or known-answer files, so these files
are the only pin: inputs come from the counter-based generator (include/b200lsap_gen.h), the
outputs are exactly what LSAP::Auction<...>::solve returned, in its own edge order.

  python tests/golden/make_golden.py            # small cases, seconds
  python tests/golden/make_golden.py --big      # + 2048^2, 4096^2 (about a minute)
  python tests/golden/make_golden.py --c4       # + 16384^2 seed 1: ~35 min of one core

Each .npz holds: rows, cols, seed, x, y (reference edge order), objective (sum of e.v in
double), variant, seconds (solve() wall time where it ran), and for the *_f32 variants the
float reference's own edges as xf, yf.
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

SMALL = [
    # (rows, cols, seed, also run Auction<float,MatrixXf>?)
    (1, 1, 1, True), (1, 5, 1, True), (2, 2, 1, True), (3, 7, 2, True), (8, 12, 1, True), (8, 12, 2, True),
    (16, 16, 1, False), (32, 48, 1, True), (32, 48, 2, True), (32, 48, 3, True), (64, 64, 3, False),
    (100, 101, 1, False), (200, 300, 1, True), (200, 300, 2, True), (200, 300, 3, True), (200, 200, 1, False),
    (256, 256, 2, False), (511, 513, 1, False), (512, 768, 1, True), (1024, 1536, 1, True),
]
BIG = [(2048, 3072, 1, True), (2048, 3072, 2, True), (2048, 2048, 1, False), (4096, 4096, 1, False),
       (8192, 12288, 1, False)]   # the last one: the multi-GPU column-sharded test (tests/test_gpu_multi.py)
C4 = [(16384, 16384, 1, False)]


def one(rows, cols, seed, with_float):
    a32 = po.generate(rows, cols, seed)                      # row-major fp32, exact grid values
    a32f = np.asfortranarray(a32)                            # Eigen's default layout
    t = time.time()
    ref = po.reference_solve(a32f, "auction_f32_as_f64")     # the oracle instantiation
    out = dict(rows=rows, cols=cols, seed=seed, variant="Auction<double> on widened fp32",
               x=ref["x"].astype(np.int32), y=ref["y"].astype(np.int32),
               objective=float(ref["v"].sum()), seconds=ref["seconds"])
    if with_float:
        rf = po.reference_solve(a32f, "auction")
        out.update(xf=rf["x"].astype(np.int32), yf=rf["y"].astype(np.int32),
                   objective_f=float(rf["v"].astype(np.float64).sum()), seconds_f=rf["seconds"])
    name = f"auction_{rows}x{cols}_s{seed}.npz"
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(f"{name}: obj {out['objective']:.9f} edges {len(out['x'])} solve {ref['seconds']:.3f}s "
          f"(total {time.time() - t:.1f}s)", flush=True)


def batch_case(batch, rows, cols, seed):
    """C3-shaped fixture: `batch` independent problems, reference looped (it has no batch API)."""
    r2c = np.empty((batch, rows), np.int32)
    r2c_f = np.empty((batch, rows), np.int32)
    for b in range(batch):
        a32 = po.generate(rows, cols, seed, problem=b)
        ref = po.reference_solve(np.asfortranarray(a32), "auction_f32_as_f64")
        r2c[b, ref["x"]] = ref["y"]
        rf = po.reference_solve(np.asfortranarray(a32), "auction")
        r2c_f[b, rf["x"]] = rf["y"]
    name = f"batch_{batch}x{rows}x{cols}_s{seed}.npz"
    np.savez_compressed(os.path.join(HERE, name), batch=batch, rows=rows, cols=cols, seed=seed, row_to_col=r2c,
                        row_to_col_f=r2c_f)
    print(f"{name}: float==double on {(r2c == r2c_f).all(axis=1).sum()}/{batch} problems", flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--big", action="store_true")
    ap.add_argument("--c4", action="store_true")
    ap.add_argument("--only-c4", action="store_true")
    args = ap.parse_args()
    po.build()
    if po.reference() is None:
        sys.exit("oracle/_ref not built: this script needs /root/reference")
    if not args.only_c4:
        for case in SMALL:
            one(*case)
        batch_case(256, 32, 48, 1)
        batch_case(64, 8, 12, 1)
        if args.big:
            for case in BIG:
                one(*case)
    if args.c4 or args.only_c4:
        for case in C4:
            one(*case)
