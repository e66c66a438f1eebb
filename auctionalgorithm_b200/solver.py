"""Host-side mirror of the reference's solver interface over the C-ABI (include/b200lsap.h).

Reference interface mirrored (relative to /root/reference/AuctionAlgorithm/):
    LSAP::Auction<Scalar, MatrixType>::solve(const MatrixType& a) -> Edges          Auction.h:44
    LSAP::AuctionMT<Scalar, MatrixType>::solve(const MatrixType& m, noOfThreads=2)  AuctionMT.h:50
    Edge{size_t x; size_t y; Scalar v;}, Edges = std::vector<Edge>                  AuctionCommon.h:30-40

Same names, same argument meaning: a dense rows x cols benefit matrix with rows <= cols and
coefficients in [0, 1] where 0 means "no edge"; the result is one edge per row.  Edges come
back sorted by row (the reference's order is the order in which columns were first won,
which a synchronous parallel auction does not reproduce; SURVEY.md §7 hard part 1).
Where the reference never returns (rows > cols, an all-zero row) B200LsapError is raised.

The C++ spelling of the same mirror is include/AuctionB200.h.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _capi
from ._capi import B200LsapError, Options, Stats

EDGE_DTYPE_F64 = np.dtype([("x", np.uint64), ("y", np.uint64), ("v", np.float64)])  # sizeof(Edge) == 24
EDGE_DTYPE_F32 = np.dtype([("x", np.uint64), ("y", np.uint64), ("v", np.float32)], align=True)


class Handle:
    """Owns a b200lsap_handle (device buffers + stream).  One per host thread."""

    def __init__(self, device: int = -1):
        self._lib = _capi.lib()
        self._h = C.c_void_p()
        rc = self._lib.b200lsap_create(C.byref(self._h), device)
        if rc != _capi.OK:
            raise B200LsapError(rc, self._lib.b200lsap_status_string(rc).decode())
        self.last_stats: Optional[dict] = None

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.b200lsap_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -------------------------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != _capi.OK:
            raise B200LsapError(rc, self._lib.b200lsap_last_error(self._h).decode())

    @staticmethod
    def _opts(eps_multiplier=None, eps_scaling=None, eps_start=None, eps_factor=None, max_rounds=None,
              driver=None, tail_rows=None, narrow_rows=None, row_cache=None, batch_bid_cap=None,
              prices_in=None) -> Options:
        o = Options()
        _capi.lib().b200lsap_options_default(C.byref(o))
        if eps_multiplier is not None:
            o.eps_multiplier = eps_multiplier
        if eps_scaling is not None:
            o.eps_scaling = int(eps_scaling)  # True / False / -1 / 2 (include/b200lsap.h)
        if eps_start is not None:
            o.eps_start = eps_start
        if eps_factor is not None:
            o.eps_factor = eps_factor
        if max_rounds is not None:
            o.max_rounds = max_rounds
        if driver is not None:
            o.driver = driver
        if row_cache is not None:
            o.row_cache = row_cache
        if tail_rows is not None:
            o.tail_rows = tail_rows
        if narrow_rows is not None:
            o.narrow_rows = narrow_rows
        if batch_bid_cap is not None:
            o.batch_bid_cap = batch_bid_cap
        if prices_in is not None:
            # warm start: a float64 numpy array (host entries; kept alive on the Options object) or a raw device
            # pointer (device entries)
            if isinstance(prices_in, np.ndarray):
                o._prices_keepalive = np.ascontiguousarray(prices_in, np.float64)
                o.prices_in = o._prices_keepalive.ctypes.data
            else:
                o.prices_in = int(prices_in)
        return o

    def set_stream(self, cuda_stream: int):
        self._check(self._lib.b200lsap_set_stream(self._h, C.c_void_p(cuda_stream)))

    def solve(self, a: np.ndarray, want_prices: bool = False, **options):
        """Host matrix in, row_to_col (int64[rows]) out.  float32 or float64, C or F order."""
        if a.ndim != 2:
            raise ValueError("need a 2-D matrix")
        if a.dtype not in (np.float32, np.float64):
            raise TypeError("use real types as coefficients (Auction.h:40): float32 or float64")
        if a.flags.c_contiguous:
            layout, ld = _capi.ROW_MAJOR, a.shape[1]
        elif a.flags.f_contiguous:
            layout, ld = _capi.COL_MAJOR, a.shape[0]
        else:
            a = np.ascontiguousarray(a)
            layout, ld = _capi.ROW_MAJOR, a.shape[1]
        rows, cols = a.shape
        r2c = np.full(max(rows, 1), -1, np.int64)
        prices = np.zeros(max(cols, 1), np.float64) if want_prices else None
        st = Stats()
        o = self._opts(**options)
        fn = self._lib.b200lsap_solve_f32 if a.dtype == np.float32 else self._lib.b200lsap_solve_f64
        rc = fn(self._h, a.ctypes.data, rows, cols, ld, layout, C.byref(o), r2c.ctypes.data,
                prices.ctypes.data if want_prices else None, C.byref(st))
        self.last_stats = st.as_dict()
        self._check(rc)
        return (r2c[:rows], prices) if want_prices else r2c[:rows]

    def solve_device(self, a_ptr: int, rows: int, cols: int, ld: int, layout: int, r2c_ptr: int,
                     prices_ptr: int = 0, dtype=np.float32, **options) -> dict:
        st = Stats()
        o = self._opts(**options)
        fn = self._lib.b200lsap_solve_device_f32 if dtype == np.float32 else self._lib.b200lsap_solve_device_f64
        rc = fn(self._h, C.c_void_p(a_ptr), rows, cols, ld, layout, C.byref(o), C.c_void_p(r2c_ptr),
                C.c_void_p(prices_ptr) if prices_ptr else None, C.byref(st))
        self.last_stats = st.as_dict()
        self._check(rc)
        return self.last_stats

    def solve_csr(self, values: np.ndarray, row_ptr: np.ndarray, col_idx: np.ndarray, rows: int, cols: int,
                  want_prices: bool = False, **options):
        """Compressed row-major sparse matrix in (Eigen's valuePtr / outerIndexPtr / innerIndexPtr, scipy's
        csr data / indptr / indices with sorted indices), row_to_col (int64[rows]) out: the reference's
        Auction<Scalar, SparseMatrix<Scalar, RowMajor>>::solve (Auction.h:166-198)."""
        values = np.ascontiguousarray(values)
        if values.dtype not in (np.float32, np.float64):
            raise TypeError("use real types as coefficients (Auction.h:40): float32 or float64")
        row_ptr = np.ascontiguousarray(row_ptr, np.int32)
        col_idx = np.ascontiguousarray(col_idx, np.int32)
        if row_ptr.shape[0] != rows + 1 or col_idx.shape[0] != values.shape[0]:
            raise ValueError("row_ptr needs rows + 1 entries and col_idx one per value")
        r2c = np.full(max(rows, 1), -1, np.int64)
        prices = np.zeros(max(cols, 1), np.float64) if want_prices else None
        st = Stats()
        o = self._opts(**options)
        fn = self._lib.b200lsap_solve_csr_f32 if values.dtype == np.float32 else self._lib.b200lsap_solve_csr_f64
        rc = fn(self._h, values.ctypes.data, row_ptr.ctypes.data, col_idx.ctypes.data, rows, cols, values.shape[0],
                C.byref(o), r2c.ctypes.data, prices.ctypes.data if want_prices else None, C.byref(st))
        self.last_stats = st.as_dict()
        self._check(rc)
        return (r2c[:rows], prices) if want_prices else r2c[:rows]

    def solve_csr_device(self, values_ptr: int, row_ptr_ptr: int, col_idx_ptr: int, rows: int, cols: int, nnz: int,
                         r2c_ptr: int, prices_ptr: int = 0, dtype=np.float32, **options) -> dict:
        st = Stats()
        o = self._opts(**options)
        fn = self._lib.b200lsap_solve_csr_device_f32 if dtype == np.float32 else self._lib.b200lsap_solve_csr_device_f64
        rc = fn(self._h, C.c_void_p(values_ptr), C.c_void_p(row_ptr_ptr), C.c_void_p(col_idx_ptr), rows, cols, nnz,
                C.byref(o), C.c_void_p(r2c_ptr), C.c_void_p(prices_ptr) if prices_ptr else None, C.byref(st))
        self.last_stats = st.as_dict()
        self._check(rc)
        return self.last_stats

    def solve_batch(self, a: np.ndarray, layout: int = _capi.ROW_MAJOR, **options) -> np.ndarray:
        """a: [batch, rows, cols] float32 (layout ROW_MAJOR) or [batch, cols, rows] memory
        order (COL_MAJOR, pass the array as [batch, rows, cols] logical shape with
        layout=COL_MAJOR and C-contiguous storage of the transposed problems)."""
        a = np.ascontiguousarray(a, dtype=np.float32)
        batch, rows, cols = a.shape
        out = np.empty((batch, rows), np.int32)  # every entry is written by the library (-1 for a failed problem)
        st = Stats()
        o = self._opts(**options)
        rc = self._lib.b200lsap_solve_batch_f32(self._h, a.ctypes.data, batch, rows, cols, layout, C.byref(o),
                                                out.ctypes.data, C.byref(st))
        self.last_stats = st.as_dict()
        self._check(rc)
        return out

    def solve_batch_device(self, a_ptr: int, batch: int, rows: int, cols: int, r2c_ptr: int,
                           layout: int = _capi.ROW_MAJOR, **options) -> dict:
        st = Stats()
        o = self._opts(**options)
        rc = self._lib.b200lsap_solve_batch_device_f32(self._h, C.c_void_p(a_ptr), batch, rows, cols, layout,
                                                       C.byref(o), C.c_void_p(r2c_ptr), C.byref(st))
        self.last_stats = st.as_dict()
        self._check(rc)
        return self.last_stats

    def generate_device(self, a_ptr: int, batch: int, rows: int, cols: int, ld: int, layout: int, seed: int,
                        problem0: int = 0):
        self._check(self._lib.b200lsap_generate_device_f32(self._h, C.c_void_p(a_ptr), batch, rows, cols, ld,
                                                           layout, seed, problem0))

    def generate_shard_device(self, a_ptr: int, rows: int, cols_local: int, ld: int, col0: int, cols_global: int,
                              seed: int, problem: int = 0):
        self._check(self._lib.b200lsap_generate_shard_device_f32(self._h, C.c_void_p(a_ptr), rows, cols_local, ld,
                                                                 col0, cols_global, seed, problem))

    # ---- column-sharded solve (one rank's side; auctionalgorithm_b200/sharded.py drives it) ----
    def shard_begin(self, a_ptr: int, rows: int, cols_local: int, ld: int, col0: int, cols_global: int,
                    **options) -> float:
        o = self._opts(**options)
        amax = C.c_double(0)
        self._check(self._lib.b200lsap_shard_begin_f32(self._h, C.c_void_p(a_ptr), rows, cols_local, ld, col0,
                                                       cols_global, C.byref(o), C.byref(amax)))
        return amax.value

    def shard_set_amax(self, amax: float):
        self._check(self._lib.b200lsap_shard_set_amax(self._h, amax))

    def shard_bid(self, rec_ptr: int) -> int:
        n = C.c_int64(0)
        self._check(self._lib.b200lsap_shard_bid(self._h, C.c_void_p(rec_ptr), C.byref(n)))
        return n.value

    def shard_resolve(self, gathered_ptr: int, n_ranks: int) -> int:
        n = C.c_int64(0)
        self._check(self._lib.b200lsap_shard_resolve(self._h, C.c_void_p(gathered_ptr), n_ranks, C.byref(n)))
        return n.value

    def shard_next_phase(self) -> bool:
        more = C.c_int(0)
        self._check(self._lib.b200lsap_shard_next_phase(self._h, C.byref(more)))
        return bool(more.value)

    def shard_finish(self, r2c_ptr: int, prices_ptr: int = 0) -> dict:
        st = Stats()
        rc = self._lib.b200lsap_shard_finish(self._h, C.c_void_p(r2c_ptr), C.c_void_p(prices_ptr) if prices_ptr else None,
                                             C.byref(st))
        self.last_stats = st.as_dict()
        self._check(rc)
        return self.last_stats

    # ---- fused column-sharded solve: the exchange runs inside the kernel (include/b200lsap.h) ----
    def shard_comm_create(self, rows_cap: int, n_ranks: int) -> bytes:
        """Allocates this rank's mailbox; returns its CUDA IPC handle (64 bytes) for ranks in other processes."""
        buf = C.create_string_buffer(_capi.IPC_HANDLE_BYTES)
        self._check(self._lib.b200lsap_shard_comm_create(self._h, rows_cap, n_ranks, buf))
        return buf.raw

    def shard_comm_ptr(self) -> int:
        return self._lib.b200lsap_shard_comm_ptr(self._h) or 0

    def shard_comm_connect(self, rank: int, n_ranks: int, ipc_handles: Optional[bytes] = None,
                           peer_ptrs: Optional[list] = None, ranks_on_this_device: int = 1):
        ptrs = None
        if peer_ptrs is not None:
            ptrs = (C.c_void_p * n_ranks)(*[C.c_void_p(p) if p else None for p in peer_ptrs])
        self._check(self._lib.b200lsap_shard_comm_connect(self._h, rank, n_ranks, ipc_handles, ptrs,
                                                          ranks_on_this_device))

    def shard_reserve(self, rows: int, cols_global: int):
        self._check(self._lib.b200lsap_shard_reserve(self._h, rows, cols_global))

    def solve_sharded_begin(self, a_ptr: int, rows: int, cols_local: int, ld: int, col0: int, cols_global: int,
                            **options):
        o = self._opts(**options)
        self._check(self._lib.b200lsap_solve_sharded_begin_f32(self._h, C.c_void_p(a_ptr), rows, cols_local, ld, col0,
                                                               cols_global, C.byref(o)))

    def solve_sharded_end(self, r2c_ptr: int, prices_ptr: int = 0) -> dict:
        st = Stats()
        rc = self._lib.b200lsap_solve_sharded_end(self._h, C.c_void_p(r2c_ptr),
                                                  C.c_void_p(prices_ptr) if prices_ptr else None, C.byref(st))
        self.last_stats = st.as_dict()
        self._check(rc)
        return self.last_stats

    def solve_sharded_device(self, a_ptr: int, rows: int, cols_local: int, ld: int, col0: int, cols_global: int,
                             r2c_ptr: int, prices_ptr: int = 0, **options) -> dict:
        self.solve_sharded_begin(a_ptr, rows, cols_local, ld, col0, cols_global, **options)
        return self.solve_sharded_end(r2c_ptr, prices_ptr)

    def debug_counters(self):
        out = (C.c_int64 * 16)()
        self._check(self._lib.b200lsap_debug_counters(self._h, out, 16))
        return list(out)

    def bid_sweep_device(self, a_ptr: int, rows: int, cols: int, ld: int, use_last_prices: bool = True):
        ms = C.c_double(0)
        nbytes = C.c_int64(0)
        self._check(self._lib.b200lsap_bid_sweep_device_f32(self._h, C.c_void_p(a_ptr), rows, cols, ld,
                                                            int(use_last_prices), C.byref(ms), C.byref(nbytes)))
        return ms.value, nbytes.value


_default_handle: Optional[Handle] = None


def default_handle() -> Handle:
    global _default_handle
    if _default_handle is None:
        _default_handle = Handle()
    return _default_handle


def edges_from_assignment(a: np.ndarray, row_to_col: np.ndarray) -> np.ndarray:
    """Edge{x, y, v = a(x, y)} per row, v taken from the caller's matrix (Auction.h:273,283)."""
    dt = EDGE_DTYPE_F64 if a.dtype == np.float64 else EDGE_DTYPE_F32
    e = np.empty(len(row_to_col), dt)
    e["x"] = np.arange(len(row_to_col), dtype=np.uint64)
    e["y"] = row_to_col.astype(np.uint64)
    e["v"] = a[np.arange(len(row_to_col)), row_to_col]
    return e


class Auction:
    """LSAP::Auction<Scalar, MatrixType> (Auction.h:19-44): `Auction.solve(a)` -> Edges."""

    @staticmethod
    def solve(a, handle: Optional[Handle] = None, **options) -> np.ndarray:
        """a: a dense numpy matrix, or a scipy.sparse matrix (the reference's SparseMatrix<Scalar, RowMajor>
        instantiation, Auction.h:166-198: stored entries are the edges)."""
        h = handle or default_handle()
        if hasattr(a, "tocsr") and not isinstance(a, np.ndarray):
            m = a.tocsr()
            if not m.has_sorted_indices:
                m = m.sorted_indices()
            rows, cols = m.shape
            r2c = h.solve_csr(m.data, m.indptr, m.indices, rows, cols, **options)
            e = np.empty(rows, EDGE_DTYPE_F64 if m.dtype == np.float64 else EDGE_DTYPE_F32)
            e["x"] = np.arange(rows, dtype=np.uint64)
            e["y"] = r2c.astype(np.uint64)
            e["v"] = np.asarray(m[np.arange(rows), r2c]).ravel()  # the stored value (Auction.h:273,283)
            return e
        r2c = h.solve(a, **options)
        return edges_from_assignment(a, r2c)


class AuctionMT:
    """LSAP::AuctionMT<Scalar, MatrixType> (AuctionMT.h:24-50).  noOfThreads is accepted for
    source compatibility and ignored: the column-interval split it controlled
    (AuctionMT.h:76-77) is done by the GPU grid.  Unlike the reference it is also correct for
    rows < cols."""

    @staticmethod
    def solve(m: np.ndarray, noOfThreads: int = 2, handle: Optional[Handle] = None, **options) -> np.ndarray:
        del noOfThreads
        return Auction.solve(m, handle=handle, **options)


def objective(edges: np.ndarray) -> float:
    """objFuncValue of the reference's driver (LSAP.cpp:12-19), summed in double."""
    return float(np.sum(edges["v"].astype(np.float64)))
