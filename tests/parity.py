"""Parity rule of SURVEY.md §8(c), shared by the GPU tests, smoke() and bench.py.

  1. sort both edge lists by row; identical -> bit-exact pass;
  2. otherwise ALL of: |objective difference| <= rows * eps_final (both summed in double),
     eps-complementary slackness of OUR assignment under OUR prices, unassigned columns priced
     no higher than any assigned column, every row assigned once, every column at most once.
"""
import glob
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_cases(max_rows=None, pattern="auction_*.npz"):
    out = []
    for p in sorted(glob.glob(os.path.join(GOLDEN_DIR, pattern))):
        g = np.load(p)
        if max_rows is None or int(g["rows"]) <= max_rows:
            out.append(p)
    return out


def golden_row_to_col(g):
    r2c = np.full(int(g["rows"]), -1, np.int64)
    r2c[g["x"]] = g["y"]
    return r2c


def is_valid_assignment(r2c, cols):
    r2c = np.asarray(r2c)
    return bool((r2c >= 0).all() and (r2c < cols).all() and len(np.unique(r2c)) == len(r2c))


def objective(a, r2c):
    return float(a[np.arange(len(r2c)), r2c].astype(np.float64).sum())


def eps_cs_violation(a, r2c, prices, chunk=2048):
    """max over rows of (best value - own value) in double; eps-CS holds iff <= eps."""
    worst = -np.inf
    rows = a.shape[0]
    p = prices.astype(np.float64)
    for r0 in range(0, rows, chunk):
        blk = a[r0:r0 + chunk].astype(np.float64)
        val = np.where(blk != 0, blk - p[None, :], -np.inf)
        own = val[np.arange(blk.shape[0]), r2c[r0:r0 + chunk]]
        worst = max(worst, float((val.max(axis=1) - own).max()))
    return worst


def check_parity(a, r2c, prices, ref_r2c, eps_final):
    """Returns 'exact' or 'within-bound'; raises AssertionError otherwise."""
    rows, cols = a.shape
    assert is_valid_assignment(r2c, cols), "not a one-to-one assignment of every row"
    if np.array_equal(np.asarray(r2c), np.asarray(ref_r2c)):
        return "exact"
    obj, ref_obj = objective(a, r2c), objective(a, ref_r2c)
    assert abs(obj - ref_obj) <= rows * eps_final, f"objective {obj} vs reference {ref_obj}"
    viol = eps_cs_violation(a, r2c, prices)
    assert viol <= eps_final * (1 + 1e-9) + 1e-15, f"eps-CS violated by {viol} > {eps_final}"
    free = np.ones(cols, bool)
    free[r2c] = False
    if free.any():
        assert prices[free].max() <= prices[~free].min(), "an unassigned column is priced above an assigned one"
    return "within-bound"


# ---- compressed row-major sparse inputs (the reference's SparseMatrix<Scalar, RowMajor> path) -----------
def make_csr(rows, cols, seed, dtype=np.float64, mod=4, thr=1):
    """Values of the counter-based generator (include/b200lsap_gen.h) at the kept positions:
    (i, j) is stored when j == i (a perfect matching always exists) or (7 i + 13 j) % mod >= thr.
    Returns (dense with zeros at missing entries, values, row_ptr, col_idx)."""
    from oracle import pyoracle as po
    a = po.generate(rows, cols, seed).astype(dtype)
    i = np.arange(rows)[:, None]
    j = np.arange(cols)[None, :]
    keep = (j == i) | ((7 * i + 13 * j) % mod >= thr)
    a = np.where(keep, a, 0).astype(dtype)
    row_ptr = np.concatenate([[0], np.cumsum(keep.sum(axis=1))]).astype(np.int32)
    col_idx = np.nonzero(keep)[1].astype(np.int32)
    values = a[keep]
    return a, values, row_ptr, col_idx
