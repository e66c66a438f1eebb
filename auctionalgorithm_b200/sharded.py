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
    """One rank of the host-driven sharded solve on its GPU, through the C-ABI (include/b200lsap.h).
    Stream contract: the handle's kernels and the exchange (NCCL on torch's streams, or torch.cat) must run on ONE
    stream, so the engine binds the handle to torch's current stream of the shard's device."""

    def __init__(self, handle, a_dev, rows: int, cols_local: int, ld: int, col0: int, cols_global: int, **options):
        import torch
        self.torch = torch
        self.h = handle
        handle.set_stream(torch.cuda.current_stream(a_dev.device).cuda_stream)
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
    except NotImplementedError:  # a backend without the flat form (a NCCL fault is a RuntimeError and must surface)
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


# ---------------------------------------------------------------------------------------------------------
# Fused form (b200lsap_solve_sharded_*): the per-round exchange runs inside one persistent kernel per rank — peer
# stores into every rank's mailbox over NVLink — and the host only wires the mailboxes once.
# ---------------------------------------------------------------------------------------------------------
def connect_fused(handle, rows_cap: int, group=None):
    """One process per GPU (torchrun): every rank creates its mailbox, the CUDA IPC handles go round with one
    all-gather, every rank maps its peers'.  Collective over `group`."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    mine = handle.shard_comm_create(rows_cap, world)
    if world == 1:
        handle.shard_comm_connect(0, 1)
        return
    dev = torch.device("cuda", torch.cuda.current_device())
    t = torch.frombuffer(bytearray(mine), dtype=torch.uint8).to(dev)
    out = torch.empty(world * len(mine), dtype=torch.uint8, device=dev)
    _all_gather_bytes(dist, out, t, group)
    handle.shard_comm_connect(rank, world, ipc_handles=bytes(out.cpu().numpy().tobytes()))
    dist.barrier(group)  # nobody starts a solve before every mailbox is mapped


def connect_fused_local(handles: Sequence, rows_cap: int, devices: Optional[Sequence[int]] = None):
    """All ranks in THIS process (one handle per rank; the same device several times, or several devices): the
    mailboxes are wired by pointer."""
    world = len(handles)
    for h in handles:
        h.shard_comm_create(rows_cap, world)
    ptrs = [h.shard_comm_ptr() for h in handles]
    devices = list(devices) if devices is not None else [0] * world
    for r, h in enumerate(handles):
        h.shard_comm_connect(r, world, peer_ptrs=ptrs, ranks_on_this_device=devices.count(devices[r]))


def solve_fused(handle, a_dev, rows: int, cols_local: int, ld: int, col0: int, cols_global: int,
                want_prices: bool = False, **options):
    """This rank's side of the fused sharded solve (collective: every rank calls it)."""
    import torch
    r2c = torch.empty(rows, dtype=torch.int32, device=a_dev.device)
    prices = torch.empty(cols_global, dtype=torch.float64, device=a_dev.device) if want_prices else None
    stats = handle.solve_sharded_device(a_dev.data_ptr(), rows, cols_local, ld, col0, cols_global, r2c.data_ptr(),
                                        prices.data_ptr() if want_prices else 0, **options)
    return r2c, prices, stats


def solve_fused_local(handles: Sequence, shards: Sequence, rows: int, parts: Sequence[Tuple[int, int]], lds: Sequence[int],
                      cols_global: int, want_prices: bool = False, **options):
    """Every rank from one host thread: all kernels are started before any is waited for."""
    import torch
    res = []
    torch.cuda.synchronize()  # the shards were written on torch's streams, the handles run on their own
    for h in handles:
        h.shard_reserve(rows, cols_global)  # no (re)allocation once the first rank's kernel is waiting for its peers
    for h, a, (c0, cl), ld in zip(handles, shards, parts, lds):
        h.solve_sharded_begin(a.data_ptr(), rows, cl, ld, c0, cols_global, **options)
    for h, a in zip(handles, shards):
        r2c = torch.empty(rows, dtype=torch.int32, device=a.device)
        prices = torch.empty(cols_global, dtype=torch.float64, device=a.device) if want_prices else None
        st = h.solve_sharded_end(r2c.data_ptr(), prices.data_ptr() if want_prices else 0)
        res.append((r2c, prices, st))
    return res


def eps_cs_violation(a_dev, r2c, prices, col0: int = 0, chunk_rows: int = 1024):
    """max over rows of (best a - p over the columns of a_dev) - (a - p of the assigned column), in double, on the
    device (checker code: torch ops).  With the full matrix (col0 = 0) eps-CS holds iff the result is <= eps."""
    import torch
    rows, cl = a_dev.shape
    p = prices[col0:col0 + cl]
    worst = torch.zeros((), dtype=torch.float64, device=a_dev.device)
    for r0 in range(0, rows, chunk_rows):
        blk = a_dev[r0:r0 + chunk_rows].to(torch.float64) - p[None, :]
        best = blk.max(dim=1).values
        j = r2c[r0:r0 + chunk_rows].long()
        inside = (j >= col0) & (j < col0 + cl)
        mine = torch.where(inside, blk.gather(1, (j - col0).clamp(0, cl - 1)[:, None])[:, 0], torch.full_like(best, float("inf")))
        worst = torch.maximum(worst, (best - mine).max())
    return float(worst.item())


def bench_leg(handle, torch, dist, rank: int, world: int, steps: int, rows: int, cols: int, seed: int):
    """bench.py's C5 leg: one rows x cols problem column-sharded over the ranks, generated per shard on the device,
    solved by the fused kernel; timed by CUDA events on the launching stream (max over ranks); checked on the device:
    every rank's replica identical (hash), eps-complementary slackness and the price post-condition on the full
    answer, and the assignment against a single-GPU solve of the whole matrix in the same run."""
    c0, cl = split_columns(cols, world)[rank]
    dev = torch.device("cuda", torch.cuda.current_device())
    t = torch.empty((rows, cl), dtype=torch.float32, device=dev)
    handle.generate_shard_device(t.data_ptr(), rows, cl, cl, c0, cols, seed)
    connect_fused(handle, rows)
    r2c, prices, st = solve_fused(handle, t, rows, cl, cl, c0, cols, want_prices=True)   # warm-up, allocates
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = []
    for _ in range(steps):
        r2c, prices, st = solve_fused(handle, t, rows, cl, cl, c0, cols, want_prices=True)
        ms.append(st["ms_solve"])
    ms_rank = sum(ms) / len(ms)
    ms_all = ms_rank
    if world > 1:
        tt = torch.tensor([ms_rank], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_all = float(tt.item())
    # ---- parity, on the device ----------------------------------------------------------------------------
    eps = 1e-5 / cols
    idx = torch.arange(rows, device=dev, dtype=torch.int64)
    h64 = int(((r2c.long() + 1) * ((idx * 2654435761) % 2147483629 + 1)).sum().item())
    hashes = [h64]
    if world > 1:
        hl = [None] * world
        dist.all_gather_object(hl, h64)
        hashes = hl
    replicas_identical = len(set(hashes)) == 1
    srt = torch.sort(r2c).values
    one_to_one = bool((r2c >= 0).all() and (srt[1:] != srt[:-1]).all() and int(srt[-1]) < cols)
    # eps-CS: every rank checks its own columns against its own (replicated) prices; the assigned column's value
    # comes from the rank that holds it
    assigned_val = torch.full((rows,), float("-inf"), dtype=torch.float64, device=dev)
    j = r2c.long()
    inside = (j >= c0) & (j < c0 + cl)
    rows_in = idx[inside]
    assigned_val[rows_in] = t[rows_in, j[inside] - c0].to(torch.float64) - prices[j[inside]]
    best_local = torch.empty(rows, dtype=torch.float64, device=dev)
    for r0 in range(0, rows, 1024):
        best_local[r0:r0 + 1024] = (t[r0:r0 + 1024].to(torch.float64) - prices[None, c0:c0 + cl]).max(dim=1).values
    if world > 1:
        dist.all_reduce(assigned_val, op=dist.ReduceOp.MAX)
        dist.all_reduce(best_local, op=dist.ReduceOp.MAX)
    cs_violation = float((best_local - assigned_val).max().item())
    owned = torch.zeros(cols, dtype=torch.bool, device=dev)
    owned[j] = True
    post = bool(prices[~owned].max().item() <= prices[owned].min().item()) if bool((~owned).any()) else True
    # ---- the whole matrix on ONE GPU in the same run (rank 0; 25.8 GB fit next to the shard) -------------------
    vs_single = None
    if world > 1:
        del t
        torch.cuda.empty_cache()
        same = torch.zeros(1, dtype=torch.int64, device=dev)
        if rank == 0:
            from .solver import Handle
            full = torch.empty((rows, cols), dtype=torch.float32, device=dev)
            handle.generate_shard_device(full.data_ptr(), rows, cols, cols, 0, cols, seed)
            h1 = Handle(torch.cuda.current_device())
            h1.shard_comm_create(rows, 1)
            h1.shard_comm_connect(0, 1)
            r1, p1, st1 = solve_fused(h1, full, rows, cols, cols, 0, cols, want_prices=True)
            same[0] = int(torch.equal(r1, r2c)) + 2 * int(torch.equal(p1, prices)) + 4 * int(st1["rounds"] == st["rounds"])
            h1.close()
            del full
        dist.broadcast(same, 0)
        vs_single = {7: "exact-vs-N1 (assignment, prices bit for bit, rounds)", 5: "assignment-and-rounds-equal",
                     1: "assignment-equal"}.get(int(same.item()), "DIFFERS-from-N1")
    parity = ("exact-vs-N1" if vs_single and vs_single.startswith("exact") else vs_single) if world > 1 else \
        "single shard (N = 1): see tests/test_gpu_sharded.py for the comparison with the dense solver"
    floor_ms = st["bids"] * cols * 4 / world / (6450.0 * 1e6)
    return {"workload": f"c5: dense {rows}x{cols} fp32 column-sharded over {world} rank(s), seed {seed}, device-generated",
            "ms_per_solve": ms_all, "ms_per_solve_this_rank": ms_rank, "rounds": st["rounds"], "bids": st["bids"],
            "us_per_round": 1e3 * ms_all / max(st["rounds"], 1),
            "bid_gbs_all_ranks": st["bids"] * cols * 4 / ms_all / 1e6, "hbm_floor_ms": floor_ms,
            "exchange": ("in-kernel: 32-byte records stored into every rank's mailbox over NVLink + flags, "
                         f"{world} ranks" if world > 1 else "none (single shard, same kernel)"),
            "exchanged_bytes_per_solve_per_rank": st["bids"] * RECORD_BYTES * (world - 1),
            "launches_per_solve": st["launches"], "parity": parity, "vs_single_gpu": vs_single,
            "replicas_identical": replicas_identical, "one_to_one_assignment": one_to_one,
            "eps_cs_violation": cs_violation, "eps": eps, "eps_cs_holds": bool(cs_violation <= eps * (1 + 1e-9) + 1e-15),
            "unassigned_prices_le_assigned": post}
