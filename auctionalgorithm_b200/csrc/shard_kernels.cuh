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

// ---------------------------------------------------------------------------------------------
// Fused form: the whole column-sharded solve in ONE persistent cooperative kernel per GPU, the
// exchange written into the kernel (north_star (5): "per-round NVLink max-bid reduction").
//
// Every rank owns a MAILBOX in its own HBM: [2 buffers][n_ranks sources][rows] records of 32 bytes, and one
// flag word per source.  Rank r's bid phase stores the record of bidding row i straight into slot
// [round & 1][r][i] of EVERY rank's mailbox — peer stores over NVLink (cudaDeviceEnablePeerAccess in one process,
// CUDA IPC mappings across the processes of a torchrun job), 32 bytes each, issued by the warp that completed the
// row while the other warps still scan: the transfer overlaps the scan row by row.  A system-scope fence, one grid
// sync, then the rank raises its flag in every mailbox (st.release.sys) and every CTA waits until the n_ranks
// flags of its own mailbox show the round (ld.acquire.sys on local memory).  Merge (AuctionMT.h:377-398: exact
// top-2 merge in interval = rank order), resolve and compaction then run on every rank's replica of the prices
// and of the assignment, identically, so nobody has to agree on anything else: the number of bidders, the phase and
// the end of the solve follow from the replicated state.  Records are addressed by ROW, so the replicas' bidder
// lists need not be in the same order (the host-driven path above needs k_shard_compact for that).
// Two buffers suffice: a rank can only write round t + 2 after it has seen every peer's flag of round t + 1, which
// a peer raises after it has finished reading round t.
// The host is not touched per round: 3 grid syncs and one flag exchange (~2 us over NVLink) per round.
// Flags count exchanges monotonically across solves (the epoch lives in the mailbox header), so nothing is reset
// between solves; a peer that never shows up ends the kernel through the spin limit with an error status instead
// of hanging the GPU.
// ---------------------------------------------------------------------------------------------
constexpr int kShardMaxRanks = 16;
constexpr int kShardHeaderBytes = 4096; // flags: rank q at byte 128 q; amax box at 2048 + 8 q; epoch at 3072

struct ShardComm
{
    unsigned char* box[kShardMaxRanks]; // box[q]: base of rank q's mailbox (header + records); box[rank] is local
    int rank, n_ranks;
    long long rows_cap;                 // records per (buffer, source) in every mailbox
    unsigned long long spin_ns;         // give up waiting for a peer after this long
};

__device__ __forceinline__ unsigned long long* shard_flag(const ShardComm& cm, int owner, int src)
{
    return reinterpret_cast<unsigned long long*>(cm.box[owner] + 128 * src);
}
__device__ __forceinline__ ShardRec* shard_slot(const ShardComm& cm, int owner, int buf, int src)
{
    return reinterpret_cast<ShardRec*>(cm.box[owner] + kShardHeaderBytes) + ((long long)buf * cm.n_ranks + src) * cm.rows_cap;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Raise this rank's flag in every mailbox (CTA 0, after a grid sync that follows every writer's system fence),
// then wait in every CTA until all sources have raised theirs in the local mailbox.  Returns false in a CTA that
// gave up (it has raised the status; the caller's next grid sync makes that visible to every CTA).
// amax != nullptr: the flag also carries this rank's max |a| (dropped into every mailbox just before the flag).
__device__ __forceinline__ bool shard_exchange(const ShardComm& cm, unsigned long long value, int* status,
                                               const volatile double* amax = nullptr)
{
    __shared__ int timed_out;
    if (threadIdx.x == 0)
        timed_out = 0;
    __syncthreads();
    if (blockIdx.x == 0 && threadIdx.x < cm.n_ranks)
    {
        if (amax)
            *reinterpret_cast<volatile double*>(cm.box[threadIdx.x] + 2048 + 8 * cm.rank) = *amax;
        st_release_sys(shard_flag(cm, threadIdx.x, cm.rank), value); // release: the amax above is visible with it
    }
    if (threadIdx.x < cm.n_ranks)
    {
        const unsigned long long* f = shard_flag(cm, cm.rank, threadIdx.x);
        const unsigned long long t0 = global_ns();
        unsigned spins = 0;
        while (ld_acquire_sys(f) < value)
        {
            if ((++spins & 1023u) == 0 && (global_ns() - t0 > cm.spin_ns || *reinterpret_cast<volatile int*>(status) != 0))
            {
                timed_out = 1;
                break;
            }
        }
    }
    __syncthreads();
    if (timed_out && threadIdx.x == 0)
        atomicCAS(status, 0, 4); // B200LSAP_ERR_CUDA: a peer did not arrive
    return timed_out == 0;
}

// How many segments a bidder's row is cut into.  split_factor (dense path) takes the largest power of two that keeps
// n * S <= warps, which leaves up to half of the machine idle (2992 bidders on 4736 warps: S = 1, 63 % busy, and the
// round lasts as long as one warp needs for a whole 393 KB row) and rounds the number of waves up when there are more
// bidders than warps (7781 bidders: 2 waves for 1.64 waves of work).  Here S is any integer and the items may take
// several waves: the S that minimises waves(n * S) * (row time / S + item overhead), with a warp streaming
// ~1.7 GB/s (four 512-byte pieces in flight per ~1.2 us of latency) and ~3 us per item for the partial result, the
// fence and the ticket.
// (Measured on 65536 x 98304, one GPU: 9.4-9.5 ms with this rule, 10.9 ms with any integer S and a per-item cost in
// microseconds, against 9.3 ms with split_factor's powers of two: the split-row path (partial result, fence, ticket)
// costs more per item than these models assume.  Kept for experiments, off by default.)
#ifndef B200LSAP_SHARD_SPLIT_MODEL
#define B200LSAP_SHARD_SPLIT_MODEL 0
#endif
constexpr bool kShardSplitModel = B200LSAP_SHARD_SPLIT_MODEL != 0;
__device__ __forceinline__ int shard_split(int n, int n_warps, int cols4, int min_seg, long long cap_items)
{
    int s_max = max(1, cols4 / max(min_seg, 128));
    s_max = (int)min((long long)s_max, max(1ll, cap_items / n));
    // S from {1, 2, 3, 4, 6, 8, 12, ...}: powers of two and their 1.5-fold; cost in whole-row scans, 2 % per item
    int best = 1;
    float best_cost = (float)((n + n_warps - 1) / n_warps) * 1.02f;
    for (int base = 2; base <= s_max && base <= 128; base *= 2)
        for (int s = base; s <= base + base / 2 && s <= s_max; s += max(1, base / 2))
        {
            const long long items = (long long)n * s;
            const float cost = (float)((items + n_warps - 1) / n_warps) * (1.0f / (float)s + 0.02f);
            if (cost < 0.97f * best_cost)
            {
                best = s;
                best_cost = cost;
            }
        }
    return best;
}

// Bid phase of a round over this rank's columns (bid_phase's scan): n_warps warps, numbered warp_id, share the n
// bidders of `list` (whole rows, or S segments per row merged by the last warp to arrive); the warp that completes a
// row stores its record into every rank's mailbox.  first_sweep: all prices are zero (round 1 of the solve), the
// scan also collects max |a| (amax_seen, per lane).
template <typename T>
__device__ __forceinline__ void shard_bid_phase(const DenseState<T>& l, const ShardComm& cm, int col0,
                                                const int* __restrict__ list, int n, int buf, bool first_sweep,
                                                float margin, int warp_id, int n_warps, int lane, float& amax_seen,
                                                long long cap_items)
{
    const int S = kShardSplitModel ? shard_split(n, n_warps, l.cols4, l.min_seg, cap_items)
                                   : split_factor(n, n_warps, l.cols4, l.min_seg);
    int seg = (l.cols4 + S - 1) / S;
    seg = (seg + 127) & ~127;
    const long long items = (long long)n * S;
    for (long long item = warp_id; item < items; item += n_warps)
    {
        const int b = (int)(item / S), sg = (int)(item - (long long)b * S);
        const int i = list[b];
        const int c0 = sg * seg, c1 = min(c0 + seg, l.cols4);
        Top2<T> t;
        if (first_sweep)
            t = scan_segment_amax(l, l.A + (long long)i * l.ld, margin, -INFINITY, c0, c1, lane, amax_seen);
        else
        {
            // the columns that were best at the row's earlier scans bound its second best from below (dense
            // path: cand_threshold): only the batches that hold the row's best columns take the exact path
            // (a property of the ROW: it holds for every segment; candidates are any true edges of the row, so
            // segments may store theirs concurrently: 8-byte stores, whichever lands is an edge)
            const float thr0 = cand_threshold(l, i, margin, lane);
            t = scan_segment<T>(l, l.A + (long long)i * l.ld, margin, thr0, c0, c1, lane);
        }
        if (t.j >= 0)
            cand_store(l, i, (t.j >> 2) & (kCand - 1), t.j, t.a);
        top2_warp_reduce_sparse(t);
        bool final_warp = S == 1;
        if (S > 1)
        {
            int ticket = 0;
            if (lane == 0)
            {
                l.partial[(long long)b * S + sg] = t;
                __threadfence();
                ticket = atomicAdd(&l.arrive[b], 1);
            }
            ticket = __shfl_sync(kFullMask, ticket, 0);
            if (ticket == S - 1)
            {
                __threadfence();
                Top2<T> m = top2_empty<T>();
                for (int k = lane; k < S; k += 32)
                    top2_merge(m, top2_load_cg(&l.partial[(long long)b * S + k]));
                top2_warp_reduce(m);
                t = m;
                final_warp = true;
                if (lane == 0)
                    l.arrive[b] = 0;
            }
        }
        if (final_warp && lane < cm.n_ranks)
        {
            ShardRec r;
            r.v = t.v;
            r.w = t.w;
            r.a = (double)t.a;
            r.j = t.j >= 0 ? t.j + col0 : -1;
            r.row = i;
            shard_slot(cm, lane, buf, cm.rank)[i] = r; // lane q writes rank q's mailbox
            // (no fence here: one system-scope fence per CTA after the phase — shard_publish — and the release of the
            // flag order every record before the flag; a fence per record cost ~40 us per round)
        }
    }
}

// End of a CTA's bid phase: its records (peer stores by many of its threads) are pushed out to system scope once, by one
// thread, after a block barrier; the grid sync and the st.release.sys of the flag that follow are cumulative over it.
__device__ __forceinline__ void shard_publish()
{
    __syncthreads();
    if (threadIdx.x == 0)
        __threadfence_system();
}

// Every rank, identically: merge the ranks' records of each bidder (rank order = column order, so top2_merge's
// "lowest column wins ties" is AuctionMT's interval order, AuctionMT.h:377-398) and place the bid on the replica.
template <typename T>
__device__ __forceinline__ void shard_merge_phase(const DenseState<T>& g, const ShardComm& cm, const int* __restrict__ list,
                                                  int n, int buf, double eps, double pmax, long long tid, long long n_threads)
{
    const ShardRec* mine = shard_slot(cm, cm.rank, buf, 0);
    for (long long b = tid; b < n; b += n_threads)
    {
        const int i = list[b];
        Top2<double> m;
        m.v = m.w = -kNoSecondBest;
        m.j = -1;
        m.a = 0.0;
        bool ok = true;
        for (int r = 0; r < cm.n_ranks; r++)
        {
            // (written by a peer during this kernel: L2 loads, never the non-coherent path)
            const double2* rp = reinterpret_cast<const double2*>(mine + r * cm.rows_cap + i);
            const double2 lo = __ldcg(rp), hi = __ldcg(rp + 1);
            Top2<double> o;
            o.v = lo.x;
            o.w = lo.y;
            o.a = hi.x;
            o.j = (int)(__double_as_longlong(hi.y) & 0xffffffffll);
            ok = ok && (int)(__double_as_longlong(hi.y) >> 32) == i;
            top2_merge(m, o);
        }
        if (!ok || m.j < 0)
        {
            g.ctrl->status = ok ? 2 : 4; // infeasible row / a record that is not this row's
            g.bid_col[i] = -1;
            continue;
        }
        const double bid = m.a - m.w + eps; // Auction.h:250
        g.bid_col[i] = m.j;
        g.bid_val[i] = bid;
        g.profits[i] = m.w - eps; // Auction.h:253
        slot_offer(&g.slots[m.j], bid, i);
        if (bid > pmax)
            atomicMax(reinterpret_cast<unsigned long long*>(&g.ctrl->pmax), (unsigned long long)__double_as_longlong(bid));
    }
}

// l: the scan view (this rank's columns; prices / pricesf offset to its first column), g: the global view (all
// columns; the replicated state).  Both share the controller block and the lists.
//
// Small rounds.  Most ROUNDS of a solve have a handful of bidders (65536 x 98304: 45 of the 64 rounds have at most
// 64) and cost three grid syncs each, ~45 us, for microseconds of work.  Bidders never increase inside an eps phase,
// so once a round's rows fit one CTA's reach (n * local columns * 4 bytes <= small_bytes) CTA 0 plays the rest of the
// phase alone — same phases, block barriers instead of grid syncs, the same exchange — while the other CTAs wait at
// one grid sync.  Every rank takes the decision from the replicated bidder count, so all switch in the same round.
template <typename T>
__global__ void __launch_bounds__(kThreads, 1) k_shard_persistent(const __grid_constant__ DenseState<T> l,
                                                                   const __grid_constant__ DenseState<T> g,
                                                                   const __grid_constant__ ShardComm cm, int col0,
                                                                   long long small_bytes)
{
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31;
    const int warp_global = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const int total_warps = gridDim.x * kWarpsPerBlock;
    const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
    const long long n_threads = (long long)gridDim.x * kThreads;
    volatile DenseCtrl* ctrl = g.ctrl;
    unsigned long long* epoch_p = reinterpret_cast<unsigned long long*>(cm.box[cm.rank] + 3072);
    const unsigned long long epoch = *epoch_p;
    unsigned long long x = 0; // exchanges of this solve so far
    int solo = ctrl->solo_seq; // single-CTA episodes so far (the same in every CTA)
    const long long cap_items = max((long long)total_warps + 64, 4ll * g.rows); // entries of l.partial (alloc_state)

    // max |a| over all shards (slack of the single-precision filter) is taken from the first sweep: every row bids
    // in round 1 and, all prices still being zero, a - (float)p is exact there, so that sweep needs no slack itself.
    // Every rank sends its own maximum along with the flag of round 1.
    double amax = 0.0;
    int cur = 0, phase = 0;
    const int n_phases = ctrl->n_phases;
    const long long max_rounds = ctrl->max_rounds;
    long long rounds = 0, bids = 0;
    double eps = ctrl->eps[0];
    while (ctrl->status == 0)
    {
        const int n = ctrl->count[cur];
        if (n == 0)
        {
            if (phase + 1 >= n_phases)
                break;
            phase++;
            eps = ctrl->eps[phase];
            grid.sync(); // everybody has read count[cur] == 0 before it is overwritten
            reset_phase(g, g.list[cur], tid, n_threads);
            if (tid == 0)
                ctrl->count[cur] = g.rows;
            grid.sync();
            continue;
        }
        if (rounds >= max_rounds)
        {
            if (tid == 0)
                ctrl->status = 3; // B200LSAP_ERR_MAX_ROUNDS (replicated state: every rank stops here)
            break;
        }
        const bool first_sweep = phase == 0 && rounds == 0; // all prices zero, every row bids
        if (!first_sweep && n <= kThreads && (long long)n * l.cols4 * (long long)sizeof(T) <= small_bytes)
        {
            // ---- the rest of this phase on CTA 0 ------------------------------------------------------------------
            solo++;
            if (blockIdx.x != 0)
                solo_wait(g.ctrl, solo); // doze instead of polling the grid barrier (dense_kernels.cuh)
            if (blockIdx.x == 0)
            {
                int m = n;
                while (m > 0 && rounds < max_rounds)
                {
                    if (threadIdx.x == 0)
                        ctrl->count[cur ^ 1] = 0;
                    const int buf = (int)(x & 1);
                    const double pmax = fmax(ctrl->pmax, amax);
                    float unused = 0.f;
                    shard_bid_phase<T>(l, cm, col0, g.list[cur], m, buf, false, filter_margin(amax, pmax), threadIdx.x >> 5,
                                       kWarpsPerBlock, lane, unused, cap_items);
                    shard_publish();
                    __syncthreads();
                    const bool ok_x = shard_exchange(cm, epoch + ++x, &g.ctrl->status);
                    if (ok_x)
                        shard_merge_phase<T>(g, cm, g.list[cur], m, buf, eps, pmax, threadIdx.x, kThreads);
                    __threadfence(); // bids, slots: performed before anybody of this CTA reads them back
                    __syncthreads();
                    if (ctrl->status != 0)
                        break;
                    resolve_phase<T>(g, g.list[cur], m, g.list[cur ^ 1], &g.ctrl->count[cur ^ 1], threadIdx.x, kThreads, lane);
                    __threadfence();
                    __syncthreads();
                    cur ^= 1;
                    rounds++;
                    bids += m;
                    m = ctrl->count[cur];
                }
                if (threadIdx.x == 0)
                {
                    if (m > 0 && ctrl->status == 0)
                        ctrl->status = 3; // the round watchdog
                    ctrl->cur = cur;
                    ctrl->rounds = rounds;
                    ctrl->bids = bids;
                    ctrl->shard_x = x;
                }
                solo_done(g.ctrl, solo);
            }
            grid.sync();
            cur = ctrl->cur;
            rounds = ctrl->rounds;
            bids = ctrl->bids;
            x = ctrl->shard_x;
            grid.sync(); // everybody has the state before CTA 0 moves on (it may overwrite count / cur next)
            continue;
        }
        if (tid == 0)
            ctrl->count[cur ^ 1] = 0;
        const int buf = (int)(x & 1);
        float amax_seen = 0.f;
#ifdef B200LSAP_TIMING
        long long tq = clock64();
#define SHTICK(k)                                                                                                  \
    do                                                                                                             \
    {                                                                                                              \
        if (tid == 0 && n <= kThreads && !first_sweep)                                                             \
        {                                                                                                          \
            const long long now_ = clock64();                                                                      \
            g.ctrl->dbg[k] += now_ - tq;                                                                           \
            tq = now_;                                                                                             \
        }                                                                                                          \
    } while (0)
#else
#define SHTICK(k)
#endif
        const double pmax = fmax(ctrl->pmax, amax);
        shard_bid_phase<T>(l, cm, col0, g.list[cur], n, buf, first_sweep, filter_margin(amax, pmax), warp_global, total_warps,
                           lane, amax_seen, cap_items);
        if (first_sweep)
        {
            // non-negative floats order like their bits; the controller keeps the maximum as a double
#pragma unroll
            for (int sft = 16; sft >= 1; sft >>= 1)
                amax_seen = fmaxf(amax_seen, __shfl_xor_sync(kFullMask, amax_seen, sft));
            if (lane == 0 && amax_seen > 0.f)
                atomicMax(reinterpret_cast<unsigned long long*>(&g.ctrl->amax),
                          (unsigned long long)__double_as_longlong((double)amax_seen));
        }
        SHTICK(0);
        shard_publish();
        SHTICK(1);
        grid.sync();
        SHTICK(2);
        if (!first_sweep && n <= kThreads)
        {
            // A round whose bidders fit one CTA's threads: CTA 0 alone waits for the ranks' flags, merges and resolves
            // (block barriers), the others go straight to the grid sync that ends the round: two grid syncs per round
            // instead of three.
            ++x;
            if (blockIdx.x == 0)
            {
                const bool ok_small = shard_exchange(cm, epoch + x, &g.ctrl->status);
                SHTICK(3);
                if (ok_small)
                    shard_merge_phase<T>(g, cm, g.list[cur], n, buf, eps, pmax, threadIdx.x, kThreads);
                __threadfence();
                __syncthreads();
                SHTICK(4);
                if (ctrl->status == 0)
                    resolve_phase<T>(g, g.list[cur], n, g.list[cur ^ 1], &g.ctrl->count[cur ^ 1], threadIdx.x, kThreads, lane);
                __threadfence();
                SHTICK(5);
            }
            grid.sync();
            SHTICK(6);
#ifdef B200LSAP_TIMING
            if (tid == 0)
                g.ctrl->dbg[7] += 1;
#endif
            if (ctrl->status != 0)
                break;
            cur ^= 1;
            rounds++;
            bids += n;
            continue;
        }
        const bool ok_x = shard_exchange(cm, epoch + ++x, &g.ctrl->status, first_sweep ? &ctrl->amax : nullptr);
        if (first_sweep)
        {
            for (int q = 0; q < cm.n_ranks; q++)
                amax = fmax(amax, __ldcg(reinterpret_cast<const double*>(cm.box[cm.rank] + 2048 + 8 * q)));
        }
        if (ok_x)
            shard_merge_phase<T>(g, cm, g.list[cur], n, buf, eps, pmax, tid, n_threads);
        grid.sync();
        if (ctrl->status != 0)
            break;
        if (first_sweep && tid == 0)
            ctrl->amax = amax; // the maximum over all ranks, for readers of the controller block
        resolve_phase<T>(g, g.list[cur], n, g.list[cur ^ 1], &g.ctrl->count[cur ^ 1], tid, n_threads, lane);
        grid.sync();
        cur ^= 1;
        rounds++;
        bids += n;
    }
    if (tid == 0)
    {
        ctrl->rounds = rounds;
        ctrl->bids = bids;
        ctrl->phase = phase;
        ctrl->cur = cur;
        ctrl->done = 1;
        *epoch_p = epoch + x;
    }
}

} // namespace b200lsap
