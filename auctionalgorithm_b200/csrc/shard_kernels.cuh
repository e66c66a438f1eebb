// shard_kernels.cuh — one big dense problem column-sharded over several GPUs (north_star (5)).
//
// This is AuctionMT's decomposition (AuctionMT.h:76-77, AuctionCommon.h:302-320: every worker
// scans its own column interval of the bidding row; AuctionMT.h:377-398: thread 0 merges the
// per-interval top-2 exactly, lowest interval first so the lowest column wins ties) with a GPU in
// the place of a worker thread and one exchange per ROUND in the place of two barrier crossings
// per BID:
//   k_shard_bid     each rank: local top-2 of every bidding row over its own columns -> records
//   (exchange)      the caller all-gathers the records (NCCL / NVLink; torch.distributed in
//                   auctionalgorithm_b200/sharded.py) — nothing else crosses GPUs
//   k_shard_merge   every rank, identically: merge the ranks' records per bidder, place the bids
//   k_resolve       (dense_kernels.cuh) winners, prices, owners, evictions — replicated state
//   k_shard_compact next round's bidders = unassigned rows in increasing row order, so that the
//                   b-th record means the same row on every rank
// Only A is sharded; prices / owners / assignment are replicated and evolve identically on every
// rank because every rank applies the same deterministic resolve to the same gathered bids.
#pragma once

#include "dense_kernels.cuh"

namespace b200lsap
{

struct __align__(16) ShardRec
{
    double v, w; // best and second best a - p over this rank's columns (-1e6: none, internal.h:17)
    double a;    // coefficient of the best, widened exactly
    int j;       // its GLOBAL column, -1: this rank has no edge for the row
    int row;     // the bidding row (checked by the merge: all ranks must agree on the order)
};

struct ShardView
{
    ShardRec* rec; // [n] this rank's records of the round
    int col0;      // first global column of the shard
};

// Bid phase of a shard: the wide-round scan of bid_phase (warp per row, or per row segment when
// bidders are few) over the local columns; instead of bidding, the warp that completes a row
// writes the local record.  s.prices / s.pricesf already point at the shard's first column.
template <typename T>
__global__ void __launch_bounds__(kThreads, 1) k_shard_bid(DenseState<T> s, ShardView sv, int cur, int n)
{
    const int lane = threadIdx.x & 31;
    const int warp_global = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const int total_warps = gridDim.x * kWarpsPerBlock;
    const double amax = s.ctrl->amax;
    const double pmax = fmax(s.ctrl->pmax, amax);
    const float margin = filter_margin(amax, pmax);
    const int* __restrict__ list = s.list[cur];
    const int S = split_factor(n, total_warps, s.cols4, s.min_seg);
    int seg = (s.cols4 + S - 1) / S;
    seg = (seg + 127) & ~127;
    const long long items = (long long)n * S;
    for (long long item = warp_global; item < items; item += total_warps)
    {
        const int b = (int)(item / S), sg = (int)(item - (long long)b * S);
        const int i = list[b];
        const int c0 = sg * seg, c1 = min(c0 + seg, s.cols4);
        Top2<T> t = scan_segment<T>(s, s.A + (long long)i * s.ld, margin, -INFINITY, c0, c1, lane);
        top2_warp_reduce_sparse(t);
        bool final_warp = S == 1;
        if (S > 1)
        {
            int ticket = 0;
            if (lane == 0)
            {
                s.partial[(long long)b * S + sg] = t;
                __threadfence();
                ticket = atomicAdd(&s.arrive[b], 1);
            }
            ticket = __shfl_sync(kFullMask, ticket, 0);
            if (ticket == S - 1)
            {
                __threadfence();
                Top2<T> m = top2_empty<T>();
                for (int k = lane; k < S; k += 32)
                    top2_merge(m, top2_load_cg(&s.partial[(long long)b * S + k]));
                top2_warp_reduce(m);
                t = m;
                final_warp = true;
                if (lane == 0)
                    s.arrive[b] = 0;
            }
        }
        if (final_warp && lane == 0)
        {
            ShardRec r;
            r.v = t.v;
            r.w = t.w;
            r.a = (double)t.a;
            r.j = t.j >= 0 ? t.j + sv.col0 : -1;
            r.row = i;
            sv.rec[b] = r;
        }
    }
}

// Every rank: merge the n_ranks records of each bidder (rank order = column order, so top2_merge's
// "lowest column wins ties" is AuctionMT's interval order) and place the bid on the replicated
// state.  gathered[r * n + b] is rank r's record of bidder b.
template <typename T>
__global__ void __launch_bounds__(256) k_shard_merge(DenseState<T> s, const ShardRec* __restrict__ gathered,
                                                     int n_ranks, int cur, int n, double eps)
{
    const double pmax = fmax(s.ctrl->pmax, s.ctrl->amax);
    if (blockIdx.x == 0 && threadIdx.x == 0)
        s.ctrl->count[cur ^ 1] = 0;
    for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < n; b += gridDim.x * blockDim.x)
    {
        const int i = s.list[cur][b];
        Top2<double> m;
        m.v = m.w = -kNoSecondBest;
        m.j = -1;
        m.a = 0.0;
        bool order_ok = true;
        for (int r = 0; r < n_ranks; r++)
        {
            const ShardRec rec = gathered[(long long)r * n + b];
            order_ok = order_ok && rec.row == i;
            Top2<double> o;
            o.v = rec.v;
            o.w = rec.w;
            o.a = rec.a;
            o.j = rec.j;
            top2_merge(m, o);
        }
        if (!order_ok)
        {
            s.ctrl->status = 4; // the ranks disagree on the bidder order: a protocol error
            s.bid_col[i] = -1;
            continue;
        }
        if (m.j < 0)
        {
            s.ctrl->status = 2; // B200LSAP_ERR_INFEASIBLE
            s.bid_col[i] = -1;
            continue;
        }
        const double bid = m.a - m.w + eps; // Auction.h:250
        s.bid_col[i] = m.j;
        s.bid_val[i] = bid;
        s.profits[i] = m.w - eps; // Auction.h:253
        slot_offer(&s.slots[m.j], bid, i);
        if (bid > pmax)
            atomicMax(reinterpret_cast<unsigned long long*>(&s.ctrl->pmax), (unsigned long long)__double_as_longlong(bid));
    }
}

// Next round's bidders: unassigned rows in increasing row order.  One CTA; thread t owns the
// contiguous rows [t * chunk, (t + 1) * chunk), counts, a block-wide exclusive scan of the 1024
// counts gives its output offset, then it writes its rows in order.
__global__ void __launch_bounds__(1024) k_shard_compact(const int* __restrict__ row2col, int rows, int* __restrict__ list,
                                                        int* count)
{
    __shared__ int warp_sum[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chunk = (rows + 1023) / 1024;
    const int r0 = min(tid * chunk, rows), r1 = min(r0 + chunk, rows);
    int mine = 0;
    for (int i = r0; i < r1; i++)
        mine += row2col[i] < 0;
    int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1)
    {
        const int o = __shfl_up_sync(kFullMask, incl, d);
        if (lane >= d)
            incl += o;
    }
    if (lane == 31)
        warp_sum[warp] = incl;
    __syncthreads();
    int before = 0, total = 0;
    for (int k = 0; k < 32; k++)
    {
        const int c = warp_sum[k];
        before += k < warp ? c : 0;
        total += c;
    }
    int out = before + incl - mine;
    for (int i = r0; i < r1; i++)
        if (row2col[i] < 0)
            list[out++] = i;
    if (tid == 0)
        *count = total;
}

} // namespace b200lsap
