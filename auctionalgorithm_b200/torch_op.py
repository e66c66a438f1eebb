"""torch.ops.b200lsap.* — the solver on device-resident tensors.

    import auctionalgorithm_b200.torch_op            # registers the ops
    cols = torch.ops.b200lsap.solve(a)               # a: [B, N, M] float32 CUDA tensor -> [B, N] int32 (column of row i)
    cols, prices = torch.ops.b200lsap.solve_dense(a, prices_in)   # a: [N, M] float32 / float64; prices_in: [M] float64 or empty

`solve` takes the batched kernel for N <= 32, M <= 64 (MHT-style data association, b200lsap_solve_batch_device_f32:
the reference's Gauss-Seidel order, exactly Auction<double>::solve per problem) and the dense solver once per problem
otherwise (b200lsap_solve_device_f32).  `solve_dense` is one problem with warm start and price export: pass the prices
a previous, similar problem returned and get this problem's back (Auction.h:58 keeps them as locals; a tracker that
re-solves nearly the same matrix every frame starts where it left off).  Everything runs on torch's current stream
of the tensor's device; nothing is copied to the host but the solver's status word.

PyTorch is plumbing here (tensors, streams): the work is libb200lsap.so, and without it the import fails loudly.
"""
from __future__ import annotations

import torch

from . import _capi
from .solver import Handle

_handles: dict[int, Handle] = {}


def _handle(device: torch.device) -> Handle:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    h = _handles.get(idx)
    if h is None:
        h = _handles[idx] = Handle(idx)
    stream = torch.cuda.current_stream(idx).cuda_stream
    if getattr(h, "_bound_stream", None) != stream:
        h.set_stream(stream)
        h._bound_stream = stream
    return h


def _check(a: torch.Tensor, dims: tuple[int, ...]):
    if not a.is_cuda:
        raise RuntimeError("b200lsap ops take CUDA tensors (there is no CPU path)")
    if a.dim() not in dims:
        raise RuntimeError(f"expected a tensor with {' or '.join(str(d) for d in dims)} dimensions, got {a.dim()}")


@torch.library.custom_op("b200lsap::solve", mutates_args=(), device_types="cuda")
def solve(a: torch.Tensor) -> torch.Tensor:
    _check(a, (3,))
    if a.dtype != torch.float32:
        raise RuntimeError("b200lsap::solve takes float32 benefit matrices")
    a = a.contiguous()
    batch, rows, cols = a.shape
    out = torch.empty((batch, rows), dtype=torch.int32, device=a.device)
    if batch == 0 or rows == 0:
        return out
    h = _handle(a.device)
    if rows <= 32 and cols <= 64:
        h.solve_batch_device(a.data_ptr(), batch, rows, cols, out.data_ptr())
    else:
        for b in range(batch):
            h.solve_device(a[b].data_ptr(), rows, cols, cols, _capi.ROW_MAJOR, out[b].data_ptr())
    return out


@solve.register_fake
def _(a):
    return a.new_empty((a.shape[0], a.shape[1]), dtype=torch.int32)


@torch.library.custom_op("b200lsap::solve_dense", mutates_args=(), device_types="cuda")
def solve_dense(a: torch.Tensor, prices_in: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    _check(a, (2,))
    if a.dtype not in (torch.float32, torch.float64):
        raise RuntimeError("b200lsap::solve_dense takes float32 or float64 benefit matrices")
    a = a.contiguous()
    rows, cols = a.shape
    out = torch.empty(rows, dtype=torch.int32, device=a.device)
    prices = torch.empty(cols, dtype=torch.float64, device=a.device)
    opts = {}
    if prices_in.numel() > 0:
        if prices_in.numel() != cols or prices_in.dtype != torch.float64 or prices_in.device != a.device:
            raise RuntimeError("prices_in: one float64 price per column, on the matrix's device (or an empty tensor)")
        keep = prices_in.contiguous()
        opts["prices_in"] = keep.data_ptr()
    import numpy as np
    h = _handle(a.device)
    h.solve_device(a.data_ptr(), rows, cols, cols, _capi.ROW_MAJOR, out.data_ptr(), prices.data_ptr(),
                   dtype=np.float32 if a.dtype == torch.float32 else np.float64, **opts)
    return out, prices


@solve_dense.register_fake
def _(a, prices_in):
    return a.new_empty((a.shape[0],), dtype=torch.int32), a.new_empty((a.shape[1],), dtype=torch.float64)
