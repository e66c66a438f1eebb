"""auctionalgorithm_b200 — B200-native (sm_100a) auction solver behind the reference's
Auction / AuctionMT solve interface.  The compute path is libb200lsap.so (CUDA + C-ABI,
include/b200lsap.h); this package is the thin host mirror.  No CPU fallback exists."""
from ._capi import (B200LsapError, COL_MAJOR, ROW_MAJOR, OK, ERR_SHAPE, ERR_INFEASIBLE, ERR_MAX_ROUNDS, ERR_CUDA,
                    ERR_ARG, ERR_NO_DEVICE, LIB_PATH, SIGNATURES, lib)
from .solver import Auction, AuctionMT, Handle, default_handle, edges_from_assignment, objective

__all__ = ["Auction", "AuctionMT", "Handle", "default_handle", "edges_from_assignment", "objective", "B200LsapError",
           "COL_MAJOR", "ROW_MAJOR", "lib", "LIB_PATH", "SIGNATURES"]
