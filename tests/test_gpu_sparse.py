"""GPU parity tests of the sparse (compressed row-major) path, b200lsap_solve_csr_* through the C-ABI,
against the reference-made golden vectors (tests/golden/csr_*.npz), the CPU oracle of
Auction<Scalar, SparseMatrix<Scalar, RowMajor>>::solve (Auction.h:166-198) and the dense CUDA path."""
import os
import subprocess

import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu

CSR_GOLDEN = parity.golden_cases(pattern="csr_*.npz")


def _case(g, dtype=np.float32):
    return parity.make_csr(int(g["rows"]), int(g["cols"]), int(g["seed"]), dtype, int(g["mod"]), int(g["thr"]))


@pytest.mark.parametrize("path", CSR_GOLDEN)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_csr_golden(handle, path, dtype):
    g = np.load(path)
    rows, cols = int(g["rows"]), int(g["cols"])
    a, values, row_ptr, col_idx = _case(g, dtype)
    r2c, prices = handle.solve_csr(values, row_ptr, col_idx, rows, cols, want_prices=True)
    verdict = parity.check_parity(a, r2c, prices, parity.golden_row_to_col(g), 1e-5 / cols)
    assert verdict == "exact", verdict
    assert parity.objective(a, r2c) == float(g["objective"])
    assert handle.last_stats["rounds"] > 0 and handle.last_stats["bids"] >= rows


@pytest.mark.parametrize("shape", [(8, 12, 1, 4, 1), (96, 96, 2, 8, 5), (150, 230, 3, 16, 12), (300, 300, 4, 16, 13),
                                   (257, 1000, 5, 32, 30)])
def test_csr_matches_oracle_and_dense_path(handle, oracle_lib, shape):
    rows, cols, seed, mod, thr = shape
    a, values, row_ptr, col_idx = parity.make_csr(rows, cols, seed, np.float64, mod, thr)
    ref = oracle_lib.oracle_solve_csr(values, row_ptr, col_idx, rows, cols)
    assert ref["status"] == oracle_lib.ORC_OK
    ref_r2c = np.full(rows, -1, np.int64)
    ref_r2c[ref["x"]] = ref["y"]
    r2c, prices = handle.solve_csr(values, row_ptr, col_idx, rows, cols, want_prices=True)
    parity.check_parity(a, r2c, prices, ref_r2c, 1e-5 / cols)
    assert np.array_equal(r2c, ref_r2c)
    # the dense path on the matrix with zeros at the missing entries runs the same rounds: identical
    # prices, bit for bit (same eps schedule: the sparse path starts eps scaling at 0.25)
    d_r2c, d_prices = handle.solve(a, want_prices=True, eps_start=0.25)
    assert np.array_equal(d_r2c, r2c) and np.array_equal(d_prices, prices)


def test_csr_larger_problem_is_optimal(handle):
    # 20000 x 20000, about 40 entries per row: checked against scipy's sparse matching and by eps-CS
    import scipy.sparse as sp
    from scipy.sparse.csgraph import min_weight_full_bipartite_matching
    n = 20000
    rng = np.random.default_rng(11)
    per_row = 40
    cols_idx = rng.integers(0, n, (n, per_row))
    cols_idx[:, 0] = rng.permutation(n)                      # a perfect matching exists
    cols_idx.sort(axis=1)
    keep = np.ones_like(cols_idx, bool)
    keep[:, 1:] = cols_idx[:, 1:] != cols_idx[:, :-1]        # drop duplicate columns inside a row
    row_ptr = np.concatenate([[0], np.cumsum(keep.sum(axis=1))]).astype(np.int32)
    col_idx = cols_idx[keep].astype(np.int32)
    values = ((rng.integers(0, 1 << 24, len(col_idx)) + 1) / float(1 << 24)).astype(np.float32)
    r2c, prices = handle.solve_csr(values, row_ptr, col_idx, n, n, want_prices=True)
    assert parity.is_valid_assignment(r2c, n)
    m = sp.csr_matrix((values.astype(np.float64), col_idx, row_ptr), shape=(n, n))
    got = np.asarray(m[np.arange(n), r2c]).ravel()
    assert (got > 0).all()                                    # every row holds a stored entry
    ri, ci = min_weight_full_bipartite_matching(m, maximize=True)
    best = float(m[ri, ci].sum())
    assert abs(float(got.sum()) - best) <= n * (1e-5 / n) + 1e-9
    # eps-CS on the stored entries under the returned prices
    eps = 1e-5 / n
    val = m.data - prices[m.indices]
    row_best = np.maximum.reduceat(val, m.indptr[:-1])
    own = got - prices[r2c]
    assert (row_best - own).max() <= eps * (1 + 1e-9) + 8 * np.spacing(max(1.0, prices.max()))


def test_csr_python_and_scipy_front_end(handle):
    import scipy.sparse as sp
    import auctionalgorithm_b200 as ab
    a, values, row_ptr, col_idx = parity.make_csr(64, 96, 9, np.float64, 8, 6)
    edges = ab.Auction.solve(sp.csr_matrix(a), handle=handle)
    r2c = handle.solve_csr(values, row_ptr, col_idx, 64, 96)
    assert np.array_equal(edges["y"].astype(np.int64), r2c)
    assert np.array_equal(edges["v"], a[np.arange(64), r2c])


def test_csr_errors_instead_of_hanging(handle):
    import auctionalgorithm_b200 as ab
    values = np.array([0.5, 0.25, 0.75], np.float64)
    with pytest.raises(ab.B200LsapError) as e:                                   # a row without an entry
        handle.solve_csr(values, np.array([0, 2, 2, 3], np.int32), np.array([0, 1, 2], np.int32), 3, 3)
    assert e.value.status == ab.ERR_INFEASIBLE
    with pytest.raises(ab.B200LsapError) as e:                                   # rows > cols
        handle.solve_csr(values, np.array([0, 1, 2, 3], np.int32), np.array([0, 1, 0], np.int32), 3, 2)
    assert e.value.status == ab.ERR_SHAPE
    # two rows that can only take column 0: no perfect matching, the auction never ends by itself
    with pytest.raises(ab.B200LsapError) as e:
        handle.solve_csr(values, np.array([0, 1, 2, 3], np.int32), np.array([0, 0, 1], np.int32), 3, 3, max_rounds=2000)
    assert e.value.status == ab.ERR_MAX_ROUNDS
    # column indices outside [0, cols) or not ascending inside a row would index device arrays out of bounds
    for bad in ([0, 3, 2], [0, -1, 2]):
        with pytest.raises(ab.B200LsapError) as e:
            handle.solve_csr(values, np.array([0, 1, 2, 3], np.int32), np.array(bad, np.int32), 3, 3)
        assert e.value.status == ab.ERR_ARG
    with pytest.raises(ab.B200LsapError) as e:
        handle.solve_csr(np.array([0.5, 0.25, 0.75, 0.1]), np.array([0, 2, 3, 4], np.int32), np.array([1, 0, 1, 2], np.int32), 3, 3)
    assert e.value.status == ab.ERR_ARG
    # the same checks on device-resident arrays (a validation kernel runs before anything is indexed)
    import torch
    out = torch.zeros(3, dtype=torch.int32, device="cuda")
    v = torch.tensor([0.5, 0.25, 0.75], dtype=torch.float64, device="cuda")
    for ptr, col, status in (([0, 1, 2, 3], [0, 7, 2], ab.ERR_ARG), ([0, 2, 2, 3], [0, 1, 2], ab.ERR_INFEASIBLE),
                             ([0, 2, 1, 3], [0, 1, 2], ab.ERR_ARG), ([1, 1, 2, 3], [0, 1, 2], ab.ERR_ARG)):
        p = torch.tensor(ptr, dtype=torch.int32, device="cuda")
        c = torch.tensor(col, dtype=torch.int32, device="cuda")
        with pytest.raises(ab.B200LsapError) as e:
            handle.solve_csr_device(v.data_ptr(), p.data_ptr(), c.data_ptr(), 3, 3, 3, out.data_ptr(), dtype=np.float64)
        assert e.value.status == status
    # the handle is still good
    r2c = handle.solve_csr(values, np.array([0, 1, 2, 3], np.int32), np.array([0, 1, 2], np.int32), 3, 3)
    assert list(r2c) == [0, 1, 2]


def test_csr_cpp_wrapper_matches(handle, tmp_path):
    # include/AuctionB200.h on a compressed row-major sparse matrix (tests/cpp/wrapper_demo.cpp "sparse")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "auctionalgorithm_b200")
    exe = str(tmp_path / "wrapper_demo")
    r = subprocess.run(["g++", "-std=c++11", "-O1", "-w", f"-I{root}/include", os.path.join(root, "tests", "cpp", "wrapper_demo.cpp"),
                        "-o", exe, f"-L{pkg}", "-lb200lsap", "-lpthread", f"-Wl,-rpath,{pkg}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rows, cols, seed = 48, 64, 3
    r = subprocess.run([exe, str(rows), str(cols), str(seed), "sparse"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().splitlines()
    got = np.array([[float(t) for t in l.split()] for l in lines[:-1]])
    a, values, row_ptr, col_idx = parity.make_csr(rows, cols, seed, np.float64, 4, 1)
    r2c = handle.solve_csr(values, row_ptr, col_idx, rows, cols)
    assert np.array_equal(got[:, 1].astype(np.int64), r2c)
    assert abs(float(lines[-1].split()[1]) - parity.objective(a, r2c)) < 1e-9
