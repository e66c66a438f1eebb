"""TEST INFRASTRUCTURE: a numpy stand-in for one rank of the column-sharded solve, with the same
six methods and the same 32-byte records as auctionalgorithm_b200.sharded.CudaShardEngine, so that
the host-side loop (solve_sharded: partitioning, all-gather layout, termination, eps phases) can run
under gloo on CPU ranks.  It restates the protocol of csrc/shard_kernels.cuh, not the product."""
import numpy as np
import torch

REC = np.dtype([("v", "<f8"), ("w", "<f8"), ("a", "<f8"), ("j", "<i4"), ("row", "<i4")])
assert REC.itemsize == 32
NO_SECOND = 1e6  # internal.h:17


def eps_schedule(rows, cols, eps_multiplier=1e-5, start=0.25, factor=5.0):
    """capi.cu make_plan: fixed eps unless the problem is square."""
    final = eps_multiplier / cols
    out = []
    if rows == cols:
        e = start
        while e > final * 1.0000001 and len(out) < 47:
            out.append(e)
            e /= factor
    return out + [final]


class NumpyShardEngine:
    def __init__(self, a_local, col0, cols_global):
        self.a = np.ascontiguousarray(a_local, dtype=np.float32)
        self.rows, self.cl = self.a.shape
        self.col0, self.cols = col0, cols_global
        self.eps = eps_schedule(self.rows, cols_global)
        self.n = 0

    def begin(self):
        self.prices = np.zeros(self.cols)
        self.owner = np.full(self.cols, -1, np.int64)
        self.r2c = np.full(self.rows, -1, np.int64)
        self.list = np.arange(self.rows)
        self.phase = 0
        self.n = self.rows
        self.rounds = self.bids = 0
        return float(np.abs(self.a).max())

    def set_amax(self, amax):
        self.amax = amax

    def new_buffer(self, nbytes):
        return torch.empty(nbytes, dtype=torch.uint8)

    def bid(self, send):
        rec = send.numpy()[: self.n * 32].view(REC)
        p = self.prices[self.col0:self.col0 + self.cl]
        for b, i in enumerate(self.list):
            row = self.a[i]
            d = np.where(row != 0, row.astype(np.float64) - p, -np.inf)
            d = np.where(d > -NO_SECOND, d, -np.inf)  # Auction.h:237: only values above the sentinel count
            j = int(np.argmax(d))
            if not np.isfinite(d[j]):
                rec[b] = (-NO_SECOND, -NO_SECOND, 0.0, -1, i)
                continue
            v = d[j]
            d[j] = -np.inf
            w = d.max()
            rec[b] = (v, w if np.isfinite(w) else -NO_SECOND, float(row[j]), self.col0 + j, i)
        return self.n

    def resolve(self, gathered, n_ranks):
        n = self.n
        recs = gathered.numpy()[: n_ranks * n * 32].view(REC).reshape(n_ranks, n)
        eps = self.eps[self.phase]
        best = {}
        for b, i in enumerate(self.list):
            v = w = -NO_SECOND
            j, a = -1, 0.0
            for r in range(n_ranks):
                o = recs[r, b]
                assert o["row"] == i
                other = o["v"] > v or (o["v"] == v and 0 <= o["j"] < j)
                lose_v = v if other else o["v"]
                keep_w = o["w"] if other else w
                w = max(lose_v, keep_w)
                if other:
                    v, j, a = o["v"], int(o["j"]), o["a"]
            assert j >= 0, "infeasible row"
            bid = a - w + eps
            cur = best.get(j)
            if cur is None or bid > cur[0] or (bid == cur[0] and i < cur[1]):
                best[j] = (bid, i)
        for j, (bid, i) in best.items():
            self.prices[j] = bid
            prev = self.owner[j]
            self.owner[j] = i
            self.r2c[i] = j
            if prev >= 0:
                self.r2c[prev] = -1
        self.rounds += 1
        self.bids += n
        self.list = np.flatnonzero(self.r2c < 0)
        self.n = len(self.list)
        return self.n

    def next_phase(self):
        if self.n != 0 or self.phase + 1 >= len(self.eps):
            return False
        self.phase += 1
        self.owner[:] = -1
        self.r2c[:] = -1
        self.list = np.arange(self.rows)
        self.n = self.rows
        return True

    def finish(self, want_prices=False):
        return (torch.from_numpy(self.r2c.astype(np.int32)), torch.from_numpy(self.prices.copy()) if want_prices else None,
                dict(rounds=self.rounds, bids=self.bids, phases=self.phase + 1))
