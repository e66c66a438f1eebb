"""CPU tests of the sparse (compressed row-major) oracle: oracle_auction_csr_* restates
Auction<Scalar, SparseMatrix<Scalar, RowMajor>>::solve (Auction.h:166-198).  Pinned against the
golden vectors the unmodified reference produced (tests/golden/csr_*.npz, make_golden_csr.py) and,
where oracle/_ref exists with the sparse entry, against the compiled reference itself."""
import numpy as np
import pytest

import parity

CSR_GOLDEN = parity.golden_cases(pattern="csr_*.npz")


def _case(g, dtype=np.float32):
    return parity.make_csr(int(g["rows"]), int(g["cols"]), int(g["seed"]), dtype, int(g["mod"]), int(g["thr"]))


@pytest.mark.parametrize("path", CSR_GOLDEN)
def test_csr_oracle_matches_golden(oracle_lib, path):
    g = np.load(path)
    _, values, row_ptr, col_idx = _case(g)
    assert len(values) == int(g["nnz"])
    out = oracle_lib.oracle_solve_csr(values.astype(np.float64), row_ptr, col_idx, int(g["rows"]), int(g["cols"]))
    assert out["status"] == oracle_lib.ORC_OK
    # identical edges in the reference's own order, identical values
    assert np.array_equal(out["x"], g["x"]) and np.array_equal(out["y"], g["y"])
    assert float(out["v"].sum()) == float(g["objective"])
    assert out["stats"]["reverse_work"] == 0


@pytest.mark.parametrize("shape", [(8, 12, 1), (100, 100, 2), (200, 300, 3), (300, 300, 4)])
def test_csr_oracle_equals_dense_oracle_on_the_densified_matrix(oracle_lib, shape):
    # a missing entry of the sparse matrix and a zero of the dense one are both "no edge"
    rows, cols, seed = shape
    a, values, row_ptr, col_idx = parity.make_csr(rows, cols, seed, np.float64, 8, 5)
    s = oracle_lib.oracle_solve_csr(values, row_ptr, col_idx, rows, cols, want_prices=True)
    d = oracle_lib.oracle_solve(np.ascontiguousarray(a), want_prices=True)
    assert np.array_equal(s["x"], d["x"]) and np.array_equal(s["y"], d["y"]) and np.array_equal(s["v"], d["v"])
    assert np.array_equal(s["prices"], d["prices"])


def test_csr_oracle_against_compiled_reference(oracle_lib):
    ref = oracle_lib.reference()
    if ref is None or not hasattr(ref, "ref_auction_csr_f64"):
        pytest.skip("oracle/_ref without the sparse entry on this machine")
    rng = np.random.default_rng(5)
    for rows, cols, density in [(8, 12, 0.5), (64, 64, 0.2), (150, 220, 0.05), (256, 256, 0.04)]:
        mask = rng.random((rows, cols)) < density
        mask[np.arange(rows), rng.permutation(cols)[:rows]] = True
        vals = (rng.integers(0, 1 << 24, (rows, cols)) + 1) / float(1 << 24)
        row_ptr = np.concatenate([[0], np.cumsum(mask.sum(axis=1))]).astype(np.int32)
        col_idx = np.nonzero(mask)[1].astype(np.int32)
        values = vals[mask]
        o = oracle_lib.oracle_solve_csr(values, row_ptr, col_idx, rows, cols)
        r = oracle_lib.reference_solve_csr(values, row_ptr, col_idx, rows, cols)
        assert np.array_equal(o["x"], r["x"]) and np.array_equal(o["y"], r["y"]) and np.array_equal(o["v"], r["v"])


def test_csr_oracle_stored_zero_is_an_edge(oracle_lib):
    # unlike the dense scan (Auction.h:134) the sparse scan skips nothing: a stored 0 can be assigned
    values = np.array([0.0, 0.5, 0.25], np.float64)
    row_ptr = np.array([0, 1, 3], np.int32)
    col_idx = np.array([0, 0, 1], np.int32)
    out = oracle_lib.oracle_solve_csr(values, row_ptr, col_idx, 2, 2)
    r2c = np.full(2, -1)
    r2c[out["x"]] = out["y"]
    assert list(r2c) == [0, 1]
