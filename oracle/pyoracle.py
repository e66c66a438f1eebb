"""TEST INFRASTRUCTURE — ctypes loaders for the CPU oracle and the compiled reference.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module; the product package never does.

  oracle()      -> liboracle.so   (oracle/auction_oracle.c: C restatement, cited there)
  reference()   -> oracle/_ref/libref_auction.so (the unmodified reference headers behind
                   oracle/ref_shim.cpp), or None when it was never built
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
COL_MAJOR, ROW_MAJOR = 0, 1
ORC_OK, ORC_BUDGET, ORC_INFEASIBLE = 0, 1, 2

_i64p = C.POINTER(C.c_int64)


class OracleStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("sweeps", "bids", "evictions", "scan_steps", "lambda_rejects",
                                         "reverse_calls", "reverse_work", "status")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


def build(verbose: bool = False) -> None:
    """make -C oracle: the C oracle always; _ref only where the reference tree exists."""
    r = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout, r.stderr)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed")


_oracle = None
_ref = None


def oracle() -> C.CDLL:
    global _oracle
    if _oracle is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        lib = C.CDLL(path)
        for name, ft in (("oracle_auction_f64", C.c_double), ("oracle_auction_f32", C.c_float)):
            fn = getattr(lib, name)
            fn.restype = C.c_int
            fn.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int64, C.c_void_p,
                           C.c_void_p, C.c_void_p, _i64p, C.c_void_p, C.c_void_p, C.POINTER(OracleStats)]
        for name in ("oracle_auction_csr_f64", "oracle_auction_csr_f32"):
            fn = getattr(lib, name)
            fn.restype = C.c_int
            fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p,
                           C.c_void_p, C.c_void_p, _i64p, C.c_void_p, C.c_void_p, C.POINTER(OracleStats)]
        for name in ("oracle_generate_f32", "oracle_generate_f64"):
            fn = getattr(lib, name)
            fn.restype = None
            fn.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_uint64, C.c_uint64]
        lib.oracle_auction_batch_f32.restype = C.c_int64
        lib.oracle_auction_batch_f32.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int,
                                                 C.c_void_p]
        _oracle = lib
    return _oracle


def reference():
    global _ref
    if _ref is None:
        path = os.path.join(_HERE, "_ref", "libref_auction.so")
        if not os.path.exists(path):
            return None
        lib = C.CDLL(path)
        for name in ("ref_auction_f64", "ref_auction_f32", "ref_auction_f32_as_f64"):
            fn = getattr(lib, name)
            fn.restype = C.c_int64
            fn.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_void_p,
                           C.c_void_p, C.POINTER(C.c_double)]
        lib.ref_auctionmt_f64.restype = C.c_int64
        lib.ref_auctionmt_f64.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int64,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
        _ref = lib
    return _ref


def _layout_of(a: np.ndarray):
    """(layout, ld) of a 2-D array without copying; copies to C order when neither fits."""
    if a.flags.c_contiguous:
        return a, ROW_MAJOR, a.shape[1]
    if a.flags.f_contiguous:
        return a, COL_MAJOR, a.shape[0]
    a = np.ascontiguousarray(a)
    return a, ROW_MAJOR, a.shape[1]


def generate(rows: int, cols: int, seed: int, problem: int = 0, dtype=np.float32, order: str = "C") -> np.ndarray:
    lib = oracle()
    a = np.empty((rows, cols), dtype=dtype, order=order)
    fn = lib.oracle_generate_f32 if dtype == np.float32 else lib.oracle_generate_f64
    fn(a.ctypes.data, rows, cols, cols if order == "C" else rows, ROW_MAJOR if order == "C" else COL_MAJOR,
       seed, problem)
    return a


def oracle_solve(a: np.ndarray, max_bids: int = -1, want_prices: bool = False):
    """Runs the C restatement in a's own dtype (float32 -> float arithmetic, float64 -> double).
    Returns dict(x, y, v, status, stats[, prices, profits]) with edges in the reference's order."""
    lib = oracle()
    a, layout, ld = _layout_of(a)
    rows, cols = a.shape
    cap = max(rows, cols)
    ex = np.empty(cap, np.int64)
    ey = np.empty(cap, np.int64)
    ev = np.empty(cap, a.dtype)
    ne = C.c_int64(0)
    st = OracleStats()
    prices = np.empty(cols, a.dtype) if want_prices else None
    profits = np.empty(rows, a.dtype) if want_prices else None
    fn = lib.oracle_auction_f64 if a.dtype == np.float64 else lib.oracle_auction_f32
    status = fn(a.ctypes.data, rows, cols, ld, layout, max_bids, ex.ctypes.data, ey.ctypes.data, ev.ctypes.data,
                C.byref(ne), prices.ctypes.data if want_prices else None,
                profits.ctypes.data if want_prices else None, C.byref(st))
    n = ne.value
    out = dict(x=ex[:n].copy(), y=ey[:n].copy(), v=ev[:n].copy(), status=status, stats=st.as_dict())
    if want_prices:
        out["prices"], out["profits"] = prices, profits
    return out


def reference_solve(a: np.ndarray, variant: str = "auction", n_threads: int = 2):
    """Runs the unmodified reference. variant: 'auction' (a's dtype), 'auction_f32_as_f64', 'mt'."""
    lib = reference()
    if lib is None:
        raise RuntimeError("oracle/_ref/libref_auction.so not built (needs /root/reference)")
    a, layout, ld = _layout_of(a)
    rows, cols = a.shape
    cap = max(rows, cols)
    ex = np.empty(cap, np.int64)
    ey = np.empty(cap, np.int64)
    sec = C.c_double(0)
    if variant == "auction":
        ev = np.empty(cap, a.dtype)
        fn = lib.ref_auction_f64 if a.dtype == np.float64 else lib.ref_auction_f32
        n = fn(a.ctypes.data, rows, cols, ld, layout, ex.ctypes.data, ey.ctypes.data, ev.ctypes.data, C.byref(sec))
    elif variant == "auction_f32_as_f64":
        assert a.dtype == np.float32
        ev = np.empty(cap, np.float64)
        n = lib.ref_auction_f32_as_f64(a.ctypes.data, rows, cols, ld, layout, ex.ctypes.data, ey.ctypes.data,
                                       ev.ctypes.data, C.byref(sec))
    elif variant == "mt":
        assert a.dtype == np.float64
        ev = np.empty(cap, np.float64)
        n = lib.ref_auctionmt_f64(a.ctypes.data, rows, cols, ld, layout, n_threads, ex.ctypes.data,
                                  ey.ctypes.data, ev.ctypes.data, C.byref(sec))
    else:
        raise ValueError(variant)
    return dict(x=ex[:n].copy(), y=ey[:n].copy(), v=ev[:n].copy(), seconds=sec.value)


def _csr_arrays(values, row_ptr, col_idx):
    values = np.ascontiguousarray(values)
    assert values.dtype in (np.float32, np.float64)
    return values, np.ascontiguousarray(row_ptr, np.int32), np.ascontiguousarray(col_idx, np.int32)


def oracle_solve_csr(values, row_ptr, col_idx, rows: int, cols: int, max_bids: int = -1, want_prices: bool = False):
    """The C restatement of Auction<S, SparseMatrix<S, RowMajor>>::solve in values' own dtype."""
    lib = oracle()
    values, row_ptr, col_idx = _csr_arrays(values, row_ptr, col_idx)
    cap = max(rows, cols)
    ex = np.empty(cap, np.int64)
    ey = np.empty(cap, np.int64)
    ev = np.empty(cap, values.dtype)
    ne = C.c_int64(0)
    st = OracleStats()
    prices = np.empty(cols, values.dtype) if want_prices else None
    profits = np.empty(rows, values.dtype) if want_prices else None
    fn = lib.oracle_auction_csr_f64 if values.dtype == np.float64 else lib.oracle_auction_csr_f32
    status = fn(values.ctypes.data, row_ptr.ctypes.data, col_idx.ctypes.data, rows, cols, max_bids, ex.ctypes.data,
                ey.ctypes.data, ev.ctypes.data, C.byref(ne), prices.ctypes.data if want_prices else None,
                profits.ctypes.data if want_prices else None, C.byref(st))
    n = ne.value
    out = dict(x=ex[:n].copy(), y=ey[:n].copy(), v=ev[:n].copy(), status=status, stats=st.as_dict())
    if want_prices:
        out["prices"], out["profits"] = prices, profits
    return out


def reference_solve_csr(values, row_ptr, col_idx, rows: int, cols: int):
    """The unmodified reference's sparse instantiation (double)."""
    lib = reference()
    if lib is None or not hasattr(lib, "ref_auction_csr_f64"):
        raise RuntimeError("oracle/_ref/libref_auction.so lacks ref_auction_csr_f64 (rebuild where /root/reference exists)")
    values, row_ptr, col_idx = _csr_arrays(np.asarray(values, np.float64), row_ptr, col_idx)
    cap = max(rows, cols)
    ex = np.empty(cap, np.int64)
    ey = np.empty(cap, np.int64)
    ev = np.empty(cap, np.float64)
    sec = C.c_double(0)
    lib.ref_auction_csr_f64.restype = C.c_int64
    lib.ref_auction_csr_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
    n = lib.ref_auction_csr_f64(values.ctypes.data, row_ptr.ctypes.data, col_idx.ctypes.data, rows, cols,
                                ex.ctypes.data, ey.ctypes.data, ev.ctypes.data, C.byref(sec))
    return dict(x=ex[:n].copy(), y=ey[:n].copy(), v=ev[:n].copy(), seconds=sec.value)


def reference_batch(a: np.ndarray, n_threads: int = 1):
    """The unmodified reference's Auction<double>::solve looped over a [batch, rows, cols] float32 array on
    n_threads threads -> (row_to_col int32 [batch, rows], seconds)."""
    lib = reference()
    if lib is None or not hasattr(lib, "ref_auction_batch_f32_as_f64"):
        raise RuntimeError("oracle/_ref/libref_auction.so lacks ref_auction_batch_f32_as_f64 (rebuild where /root/reference exists)")
    a = np.ascontiguousarray(a, dtype=np.float32)
    batch, rows, cols = a.shape
    out = np.full((batch, rows), -1, np.int32)
    sec = C.c_double(0)
    lib.ref_auction_batch_f32_as_f64.restype = C.c_int64
    lib.ref_auction_batch_f32_as_f64.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                                 C.POINTER(C.c_double)]
    lib.ref_auction_batch_f32_as_f64(a.ctypes.data, batch, rows, cols, n_threads, out.ctypes.data, C.byref(sec))
    return out, sec.value


def oracle_batch(a: np.ndarray, use_double: bool = True, n_threads: int = 1):
    """a: [batch, rows, cols] float32 C-contiguous -> (row_to_col int32 [batch, rows], failed)."""
    lib = oracle()
    a = np.ascontiguousarray(a, dtype=np.float32)
    batch, rows, cols = a.shape
    out = np.empty((batch, rows), np.int32)
    failed = lib.oracle_auction_batch_f32(a.ctypes.data, batch, rows, cols, int(use_double), n_threads,
                                          out.ctypes.data)
    return out, int(failed)
