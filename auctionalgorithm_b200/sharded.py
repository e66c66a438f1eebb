"""Column-sharded solve of one big dense problem over several GPUs (north_star (5)).

This is the host side of AuctionMT's decomposition (AuctionMT.h:76-77: `splitToIntervals(n, cols)`
gives every worker a column interval; AuctionMT.h:366-398: per bid, every worker scans its interval
and thread 0 merges the top-2 in interval order) with one GPU per interval and one exchange per
ROUND: every rank scans its own columns for all bidding rows, the ranks all-gather 32-byte records
(NCCL over NVLink through torch.distributed; nothing else crosses GPUs), and every rank applies the
same deterministic merge + resolve to its replica of the prices and of the assignment.

    engine   one rank's kernels: CudaShardEngine (the C-ABI b200lsap_shard_* calls), or any object
             with the same six methods (tests/ run the loop under gloo with a CPU stand-in)
    exchange all-gather of a byte tensor: torch.distributed, or a local concatenation when several
             shards are emulated in one process (tests on a single GPU)
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

RECORD_BYTES = 32  # struct ShardRec {double v, w, a; int32 j, row;}  (csrc/shard_kernels.cuh)


def split_columns(cols: int, n_ranks: int, align: int = 128) -> List[Tuple[int, int]]:
    """(col0, cols_local) per rank: AuctionCommon.h:302-320 (cumulative interval ends, the remainder
    spread over the first intervals), with the boundaries rounded to `align` columns so that every
    shard starts on a 512-byte boundary.  Trailing ranks may get fewer (never zero) columns."""
    if n_ranks < 1 or cols < n_ranks:
        raise ValueError("need 1 <= n_ranks <= cols")
    blocks = -(-cols // align)
    if blocks < n_ranks:
        align, blocks = 4, -(-cols // 4)
        if blocks < n_ranks:
            raise ValueError(f"{cols} columns cannot be split over {n_ranks} ranks in multiples of 4")
    base, rem = divmod(blocks, n_ranks)
    out, c = [], 0
    for r in range(n_ranks):
        nb = base + (1 if r < rem else 0)
        c1 = min(cols, c + nb * align)
        out.append((c, c1 - c))
        c = c1
    assert c == cols and all(n > 0 for _, n in out)
    return out


class CudaShardEngine:
    """One rank of the sharded solve on its GPU, through the C-ABI (include/b200lsap.h)."""

    def __init__(self, handle, a_dev, rows: int, cols_local: int, ld: int, col0: int, cols_global: int, **options):
        import torch
        self.torch = torch
        self.h = handle
        self.a = a_dev  # keeps the shard alive
        self.shape = (rows, cols_local, ld, col0, cols_global)
        self.options = options
        self.device = a_dev.device
        self.n = 0

    def begin(self) -> float:
        rows, cols_local, ld, col0, cols_global = self.shape
        amax = self.h.shard_begin(self.a.data_ptr(), rows, cols_local, ld, col0, cols_global, **self.options)
        self.n = rows
        return amax

    def set_amax(self, amax: float):
        self.h.shard_set_amax(amax)

    def new_buffer(self, nbytes: int):
        return self.torch.empty(nbytes, dtype=self.torch.uint8, device=self.device)

    def bid(self, send) -> int:
        return self.h.shard_bid(send.data_ptr())

    def resolve(self, gathered, n_ranks: int) -> int:
        self.n = self.h.shard_resolve(gathered.data_ptr(), n_ranks)
        return self.n

    def next_phase(self) -> bool:
        more = self.h.shard_next_phase()
        if more:
            self.n = self.shape[0]
        return more

    def finish(self, want_prices: bool = False):
        rows, _, _, _, cols_global = self.shape
        r2c = self.torch.empty(rows, dtype=self.torch.int32, device=self.device)
        prices = self.torch.empty(cols_global, dtype=self.torch.float64, device=self.device) if want_prices else None
        stats = self.h.shard_finish(r2c.data_ptr(), prices.data_ptr() if want_prices else 0)
        return r2c, prices, stats


def _all_gather_bytes(dist, out, inp, group):
    try:
        dist.all_gather_into_tensor(out, inp, group=group)
    except (RuntimeError, NotImplementedError):  # a backend without the flat form
        world = dist.get_world_size(group)
        parts = list(out.chunk(world))
        dist.all_gather(parts, inp, group=group)


def solve_sharded(engine, rows: int, group=None, want_prices: bool = False):
    """Runs this rank's side of the sharded auction over torch.distributed (world = the ranks of
    `group`, in increasing col0 order).  Returns (row_to_col, prices, stats) of this rank's
    replica; every rank ends with the same assignment."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    amax = torch.tensor([engine.begin()], dtype=torch.float64)
    send = engine.new_buffer(rows * RECORD_BYTES)
    recv = engine.new_buffer(world * rows * RECORD_BYTES)
    amax = amax.to(send.device)
    dist.all_reduce(amax, op=dist.ReduceOp.MAX, group=group)
    engine.set_amax(float(amax.item()))
    exchanged = 0
    while True:
        n = engine.n
        while n > 0:
            got = engine.bid(send)
            assert got == n
            nbytes = n * RECORD_BYTES
            _all_gather_bytes(dist, recv[:world * nbytes], send[:nbytes], group)
            exchanged += world * nbytes
            n = engine.resolve(recv, world)
        if not engine.next_phase():
            break
    r2c, prices, stats = engine.finish(want_prices)
    stats = dict(stats, exchanged_bytes=exchanged)
    return r2c, prices, stats


def solve_sharded_emulated(engines: Sequence, rows: int, want_prices: bool = False):
    """All shards in ONE process (e.g. several handles on one GPU): the all-gather becomes a
    concatenation.  Used by the single-GPU tests; same engine calls in the same order."""
    import torch
    world = len(engines)
    amax = max(e.begin() for e in engines)
    for e in engines:
        e.set_amax(amax)
    sends = [e.new_buffer(rows * RECORD_BYTES) for e in engines]
    while True:
        n = engines[0].n
        while n > 0:
            for e, sbuf in zip(engines, sends):
                assert e.bid(sbuf) == n
            gathered = torch.cat([sbuf[:n * RECORD_BYTES] for sbuf in sends])
            nexts = [e.resolve(gathered, world) for e in engines]
            assert len(set(nexts)) == 1, f"replicas diverged: {nexts}"
            n = nexts[0]
        more = [e.next_phase() for e in engines]
        assert len(set(more)) == 1
        if not more[0]:
            break
    return [e.finish(want_prices) for e in engines]
