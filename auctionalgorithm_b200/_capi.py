"""ctypes binding of include/b200lsap.h.  Loading fails loudly when libb200lsap.so is missing:
there is no CPU path behind this package."""
from __future__ import annotations

import ctypes as C
import os

PKG = os.path.dirname(os.path.abspath(__file__))
# B200LSAP_LIB: another build of the same library (e.g. the -DB200LSAP_TIMING developer build)
LIB_PATH = os.environ.get("B200LSAP_LIB") or os.path.join(PKG, "libb200lsap.so")

COL_MAJOR, ROW_MAJOR = 0, 1
IPC_HANDLE_BYTES = 64
NOALLOC_MAX_ELEMS = 2097152
OK, ERR_SHAPE, ERR_INFEASIBLE, ERR_MAX_ROUNDS, ERR_CUDA, ERR_ARG, ERR_NO_DEVICE = range(7)


class Options(C.Structure):
    _fields_ = [("eps_multiplier", C.c_double), ("eps_scaling", C.c_int32), ("eps_start", C.c_double),
                ("eps_factor", C.c_double), ("max_rounds", C.c_int64), ("driver", C.c_int32),
                ("tail_rows", C.c_int32), ("narrow_rows", C.c_int32), ("row_cache", C.c_int32),
                ("batch_bid_cap", C.c_int32), ("prices_in", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("rounds", C.c_int64), ("bids", C.c_int64), ("phases", C.c_int64), ("bid_bytes", C.c_int64),
                ("eps_final", C.c_double), ("ms_upload", C.c_double), ("ms_solve", C.c_double),
                ("ms_download", C.c_double), ("launches", C.c_int64), ("tail_bids", C.c_int64), ("tail_runs", C.c_int64),
                ("narrow_rounds", C.c_int64), ("cache_hits", C.c_int64), ("cache_misses", C.c_int64),
                ("eps_first", C.c_double), ("batch_takeovers", C.c_int64), ("batch_max_bids", C.c_int64),
                ("warm_fallback", C.c_int64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


# every symbol include/b200lsap.h declares: (name, restype, argtypes)
_H = C.c_void_p
_P = C.c_void_p
_i64 = C.c_int64
SIGNATURES = {
    "b200lsap_create": (C.c_int, [C.POINTER(_H), C.c_int]),
    "b200lsap_destroy": (C.c_int, [_H]),
    "b200lsap_set_stream": (C.c_int, [_H, _P]),
    "b200lsap_last_error": (C.c_char_p, [_H]),
    "b200lsap_status_string": (C.c_char_p, [C.c_int]),
    "b200lsap_options_default": (None, [C.POINTER(Options)]),
    "b200lsap_solve_f32": (C.c_int, [_H, _P, _i64, _i64, _i64, C.c_int, C.POINTER(Options), _P, _P, C.POINTER(Stats)]),
    "b200lsap_solve_f64": (C.c_int, [_H, _P, _i64, _i64, _i64, C.c_int, C.POINTER(Options), _P, _P, C.POINTER(Stats)]),
    "b200lsap_solve_device_f32": (C.c_int, [_H, _P, _i64, _i64, _i64, C.c_int, C.POINTER(Options), _P, _P,
                                            C.POINTER(Stats)]),
    "b200lsap_solve_device_f64": (C.c_int, [_H, _P, _i64, _i64, _i64, C.c_int, C.POINTER(Options), _P, _P,
                                            C.POINTER(Stats)]),
    "b200lsap_solve_csr_f32": (C.c_int, [_H, _P, _P, _P, _i64, _i64, _i64, C.POINTER(Options), _P, _P, C.POINTER(Stats)]),
    "b200lsap_solve_csr_f64": (C.c_int, [_H, _P, _P, _P, _i64, _i64, _i64, C.POINTER(Options), _P, _P, C.POINTER(Stats)]),
    "b200lsap_solve_csr_device_f32": (C.c_int, [_H, _P, _P, _P, _i64, _i64, _i64, C.POINTER(Options), _P, _P,
                                                C.POINTER(Stats)]),
    "b200lsap_solve_csr_device_f64": (C.c_int, [_H, _P, _P, _P, _i64, _i64, _i64, C.POINTER(Options), _P, _P,
                                                C.POINTER(Stats)]),
    "b200lsap_solve_batch_f32": (C.c_int, [_H, _P, _i64, C.c_int32, C.c_int32, C.c_int, C.POINTER(Options), _P,
                                           C.POINTER(Stats)]),
    "b200lsap_solve_batch_device_f32": (C.c_int, [_H, _P, _i64, C.c_int32, C.c_int32, C.c_int, C.POINTER(Options), _P,
                                                  C.POINTER(Stats)]),
    "b200lsap_generate_device_f32": (C.c_int, [_H, _P, _i64, _i64, _i64, _i64, C.c_int, C.c_uint64, C.c_uint64]),
    "b200lsap_generate_shard_device_f32": (C.c_int, [_H, _P, _i64, _i64, _i64, _i64, _i64, C.c_uint64, C.c_uint64]),
    "b200lsap_bid_sweep_device_f32": (C.c_int, [_H, _P, _i64, _i64, _i64, C.c_int, C.POINTER(C.c_double),
                                                C.POINTER(_i64)]),
    "b200lsap_shard_record_bytes": (_i64, []),
    "b200lsap_shard_begin_f32": (C.c_int, [_H, _P, _i64, _i64, _i64, _i64, _i64, C.POINTER(Options),
                                           C.POINTER(C.c_double)]),
    "b200lsap_shard_set_amax": (C.c_int, [_H, C.c_double]),
    "b200lsap_shard_bid": (C.c_int, [_H, _P, C.POINTER(_i64)]),
    "b200lsap_shard_resolve": (C.c_int, [_H, _P, C.c_int, C.POINTER(_i64)]),
    "b200lsap_shard_next_phase": (C.c_int, [_H, C.POINTER(C.c_int)]),
    "b200lsap_shard_finish": (C.c_int, [_H, _P, _P, C.POINTER(Stats)]),
    "b200lsap_shard_comm_create": (C.c_int, [_H, _i64, C.c_int, _P]),
    "b200lsap_shard_comm_ptr": (C.c_void_p, [_H]),
    "b200lsap_shard_comm_connect": (C.c_int, [_H, C.c_int, C.c_int, _P, _P, C.c_int]),
    "b200lsap_shard_reserve": (C.c_int, [_H, _i64, _i64]),
    "b200lsap_solve_sharded_begin_f32": (C.c_int, [_H, _P, _i64, _i64, _i64, _i64, _i64, C.POINTER(Options)]),
    "b200lsap_solve_sharded_end": (C.c_int, [_H, _P, _P, C.POINTER(Stats)]),
    "b200lsap_solve_sharded_device_f32": (C.c_int, [_H, _P, _i64, _i64, _i64, _i64, _i64, C.POINTER(Options), _P, _P,
                                                    C.POINTER(Stats)]),
    "b200lsap_create_multi": (C.c_int, [C.POINTER(_H), C.POINTER(C.c_int), C.c_int]),
    "b200lsap_destroy_multi": (C.c_int, [_H]),
    "b200lsap_multi_last_error": (C.c_char_p, [_H]),
    "b200lsap_multi_devices": (C.c_int, [_H]),
    "b200lsap_multi_solve_f32": (C.c_int, [_H, _P, _i64, _i64, _i64, C.c_int, C.POINTER(Options), _P, _P,
                                           C.POINTER(Stats)]),
    "b200lsap_multi_solve_batch_f32": (C.c_int, [_H, _P, _i64, C.c_int32, C.c_int32, C.c_int, C.POINTER(Options), _P,
                                                 C.POINTER(Stats)]),
    "b200lsap_solve_batch_noalloc_f32": (C.c_int, [C.c_int, _P, _P, _i64, C.c_int32, C.c_int32, C.c_int, C.POINTER(Options),
                                                   _P, _P]),
    "b200lsap_noalloc_last_error": (C.c_char_p, []),
    "b200lsap_debug_counters": (C.c_int, [_H, C.POINTER(_i64), C.c_int]),
    "b200lsap_host_alloc": (C.c_void_p, [_i64]),
    "b200lsap_host_free": (None, [_P]),
}

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m auctionalgorithm_b200.build` "
                "(nvcc, sm_100a). auctionalgorithm_b200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError here = header and library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class B200LsapError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"b200lsap status {status}: {message}")
        self.status = status
