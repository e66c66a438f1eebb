"""Golden vectors of the reference's sparse instantiation, made by running the UNMODIFIED reference
(oracle/_ref/libref_auction.so: Auction<double, Eigen::SparseMatrix<double, RowMajor>>::solve,
Auction.h:166-198; the driver's call at LSAP.cpp:51) on compressed row-major matrices from
tests/parity.make_csr (generator values at a fixed sparsity pattern).

  python tests/golden/make_golden_csr.py

Each csr_*.npz holds rows, cols, seed, mod, thr (the pattern), x, y (reference edge order) and
objective; fp32 cases ran Auction<double> on the exactly widened values.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import pyoracle as po  # noqa: E402
import parity  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [  # rows, cols, seed, mod, thr
    (1, 1, 1, 4, 1), (8, 12, 1, 4, 1), (32, 48, 2, 4, 1), (64, 64, 3, 4, 1), (200, 300, 1, 16, 13),
    (200, 200, 1, 16, 13), (512, 512, 1, 16, 12), (1024, 1536, 2, 32, 29), (2048, 2048, 1, 64, 60),
]


def main():
    for rows, cols, seed, mod, thr in CASES:
        _, values, row_ptr, col_idx = parity.make_csr(rows, cols, seed, np.float32, mod, thr)
        ref = po.reference_solve_csr(values.astype(np.float64), row_ptr, col_idx, rows, cols)
        name = f"csr_{rows}x{cols}_s{seed}_m{mod}t{thr}.npz"
        np.savez_compressed(os.path.join(HERE, name), rows=rows, cols=cols, seed=seed, mod=mod, thr=thr,
                            nnz=len(values), x=ref["x"].astype(np.int32), y=ref["y"].astype(np.int32),
                            objective=float(ref["v"].sum()), seconds=ref["seconds"],
                            variant="Auction<double, SparseMatrix<double, RowMajor>> on widened fp32")
        print(f"{name}: nnz {len(values)} obj {ref['v'].sum():.9f} edges {len(ref['x'])} solve {ref['seconds']:.3f}s",
              flush=True)


if __name__ == "__main__":
    main()
