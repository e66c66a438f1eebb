// dense_kernels.cuh — the auction round for one dense problem on one GPU.
//
// A round is the synchronous (Jacobi) form of the reference's Gauss-Seidel sweep:
//   bid phase      every unassigned row scans its row of A against the same prices
//                  (findForward, Auction.h:124-152; forwardGS, AuctionCommon.h:86-129) and offers
//                  bid = a - w + eps (Auction.h:250) to its best column;
//   resolve phase  per column the highest bid wins (lowest row among equal bids), the price
//                  becomes the bid, the previous owner is evicted (Auction.h:257-285, with
//                  owner[col] / row2col[row] arrays in place of the O(|E|) edge scan) and the
//                  next round's list of unassigned rows is compacted.
// The controller (Auction.h:44-109: run forward sweeps until nobody is left, eps fixed or
// scaled) runs on the device in k_auction_persistent; cooperative grid syncs stand where
// AuctionMT crosses its ThreadBarrier twice per bid (AuctionMT.h:366-368).
#pragma once

#include <cooperative_groups.h>

#include "dense_state.cuh"

namespace b200lsap
{
namespace cg = cooperative_groups;

constexpr int kThreads = 1024;        // one CTA per SM, 32 warps
constexpr int kWarpsPerBlock = kThreads / 32;

// ---------------------------------------------------------------------------------------------
// Warp scan of columns [c0, c1) of one row (c0, c1 multiples of 4, row 16-byte aligned):
// each lane streams 128-bit pieces 512 bytes apart, four in flight, then a shuffle reduction.
// This is the exact form: every element goes through the double-precision top-2 update.
// ---------------------------------------------------------------------------------------------
template <typename T, bool kDeep = false>
__device__ __forceinline__ Top2<T> scan_segment_exact(const T* __restrict__ row, const double* prices, int c0, int c1,
                                                      int lane)
{
    using V4 = typename Vec4<T>::type;
    Top2<T> t = top2_empty<T, kDeep>();
    constexpr int U = 4;
    for (int c = c0 + lane * 4; c < c1; c += 128 * U)
    {
        V4 a[U];
        double2 p[U][2];
#pragma unroll
        for (int u = 0; u < U; u++)
        {
            const int cc = c + u * 128;
            if (cc < c1)
            {
                a[u] = ld_stream4(row + cc);
                p[u][0] = *reinterpret_cast<const double2*>(prices + cc);
                p[u][1] = *reinterpret_cast<const double2*>(prices + cc + 2);
            }
            else
            {
                a[u].x = a[u].y = a[u].z = a[u].w = T(0);
                p[u][0] = p[u][1] = make_double2(0., 0.);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++)
        {
            const int cc = c + u * 128;
            top2_push(t, a[u].x, p[u][0].x, cc);
            top2_push(t, a[u].y, p[u][0].y, cc + 1);
            top2_push(t, a[u].z, p[u][1].x, cc + 2);
            top2_push(t, a[u].w, p[u][1].y, cc + 3);
        }
    }
    top2_warp_reduce(t);
    return t;
}

// ---------------------------------------------------------------------------------------------
// fp32 rows: the same scan behind a single-precision filter.  f = a - (float)p is within
// 2^-23 (amax + pmax) of the exact double a - p, and an element can only change the row's top-2 if
// its exact value exceeds the row's current second best.  `thr` is a warp-uniform lower bound of
// that second best, lowered by `margin` >= 2^-21 (amax + pmax + 1): elements with f <= thr are
// provably irrelevant and cost one FADD and one FMNMX; the few that pass (about 2 ln(cols) per
// row) take the exact double path of top2_push.  The result is bit-identical to
// scan_segment_exact; only the instruction count and the L1 price traffic (4 B instead of 8 B
// per element) change.  Lane l visits its columns in increasing order, as the exact scan does.
// ---------------------------------------------------------------------------------------------
template <bool kDeep = false>
__device__ __forceinline__ Top2<float> scan_segment_filtered(const float* __restrict__ row, const double* prices,
                                                             const float* pricesf, float margin, int c0, int c1,
                                                             int lane)
{
    Top2<float> t = top2_empty<float, kDeep>();
    float thr = -CUDART_INF_F;
    constexpr int U = 4;
    for (int base = c0; base < c1; base += 128 * U) // warp-uniform trip count
    {
        float4 a[U];
        float m = -CUDART_INF_F;
#pragma unroll
        for (int u = 0; u < U; u++)
        {
            const int cc = base + u * 128 + lane * 4;
            a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (cc < c1)
                a[u] = ld_stream4(row + cc);
        }
#pragma unroll
        for (int u = 0; u < U; u++)
        {
            const int cc = base + u * 128 + lane * 4;
            if (cc < c1)
            {
                const float4 q = *reinterpret_cast<const float4*>(pricesf + cc);
                m = fmaxf(m, fmaxf(fmaxf(a[u].x - q.x, a[u].y - q.y), fmaxf(a[u].z - q.z, a[u].w - q.w)));
            }
        }
        if (__any_sync(kFullMask, m > thr))
        {
            // rare: redo the subtraction (prices hit L1) and send what passes through the exact update
#pragma unroll
            for (int u = 0; u < U; u++)
            {
                const int cc = base + u * 128 + lane * 4;
                if (cc < c1)
                {
                    const float4 q = *reinterpret_cast<const float4*>(pricesf + cc);
                    if (a[u].x - q.x > thr)
                        top2_push(t, a[u].x, prices[cc], cc);
                    if (a[u].y - q.y > thr)
                        top2_push(t, a[u].y, prices[cc + 1], cc + 1);
                    if (a[u].z - q.z > thr)
                        top2_push(t, a[u].z, prices[cc + 2], cc + 2);
                    if (a[u].w - q.w > thr)
                        top2_push(t, a[u].w, prices[cc + 3], cc + 3);
                }
            }
            // second largest of all lanes' (best, second best), rounded down: a lower bound of the
            // row's second best so far
            float hi = __double2float_rd(t.v), lo = __double2float_rd(t.w);
#pragma unroll
            for (int s = 16; s >= 1; s >>= 1)
            {
                const float ohi = __shfl_xor_sync(kFullMask, hi, s);
                const float olo = __shfl_xor_sync(kFullMask, lo, s);
                lo = fmaxf(fminf(hi, ohi), fmaxf(lo, olo));
                hi = fmaxf(hi, ohi);
            }
            thr = fmaxf(thr, __fsub_rd(lo, margin));
        }
    }
    top2_warp_reduce(t);
    return t;
}

template <typename T, bool kDeep = false>
__device__ __forceinline__ Top2<T> scan_segment(const DenseState<T>& s, const T* __restrict__ row, float margin, int c0,
                                                int c1, int lane)
{
    if constexpr (sizeof(T) == 4)
        return scan_segment_filtered<kDeep>(row, s.prices, s.pricesf, margin, c0, c1, lane);
    else
        return scan_segment_exact<T, kDeep>(row, s.prices, c0, c1, lane);
}

// All lanes of the deciding warp.  A row whose every a - p lies at or below -1e6 (possible once
// a single-edge row has paid the 1e6 sentinel and eps scaling makes it bid again) is re-scanned
// without the sentinel; with no second candidate it raises its column by exactly eps.
template <typename T>
__device__ __forceinline__ Top2<T> rescue_scan(const DenseState<T>& s, int i, const Top2<T>& t, int lane)
{
    if (t.j >= 0)
        return t;
    Top2<T> d = scan_segment_exact<T, true>(s.A + (long long)i * s.ld, s.prices, 0, s.cols4, lane);
    if (d.j >= 0 && d.w == -CUDART_INF)
        d.w = fmin(-kNoSecondBest, d.v);
    return d;
}

// Lane 0 of the deciding warp: record and offer the bid of row i.
template <typename T>
__device__ __forceinline__ void submit_bid(const DenseState<T>& s, int i, const Top2<T>& t, double eps, double pmax_seen)
{
    if (t.j < 0)
    {
        s.ctrl->status = 2; // B200LSAP_ERR_INFEASIBLE: the reference would never return
        s.bid_col[i] = -1;
        return;
    }
    const double bid = (double)t.a - t.w + eps; // Auction.h:250, same order of operations
    s.bid_col[i] = t.j;
    s.bid_val[i] = bid;
    s.profits[i] = t.w - eps;                   // Auction.h:253
    if (!s.dry)
    {
        slot_offer(&s.slots[t.j], bid, i);
        if (bid > pmax_seen) // rare: prices seldom exceed the largest coefficient
            atomicMax(reinterpret_cast<unsigned long long*>(&s.ctrl->pmax), (unsigned long long)__double_as_longlong(bid));
    }
}

// How many warps share one row this round: 1 while there are at least as many bidders as warps,
// otherwise the largest power of two that keeps every warp busy and segments >= min_seg.
__device__ __forceinline__ int split_factor(int n, int total_warps, int cols4, int min_seg)
{
    if (n >= total_warps)
        return 1;
    int s = 1;
    while (s * 2 * n <= total_warps && (cols4 + s * 2 - 1) / (s * 2) >= min_seg)
        s *= 2;
    return s;
}

template <typename T>
__device__ __forceinline__ void bid_phase(const DenseState<T>& s, const int* __restrict__ list, int n, double eps,
                                          double amax, double pmax, int warp_global, int total_warps, int lane)
{
    const float margin = filter_margin(amax, pmax);
    const int S = split_factor(n, total_warps, s.cols4, s.min_seg);
    int seg = (s.cols4 + S - 1) / S;
    seg = (seg + 127) & ~127; // segments start on 512-byte boundaries
    const long long items = (long long)n * S;
    for (long long item = warp_global; item < items; item += total_warps)
    {
        const int b = (int)(item / S), sg = (int)(item - (long long)b * S);
        const int i = list[b];
        const int c0 = sg * seg;
        const int c1 = min(c0 + seg, s.cols4);
        Top2<T> t = scan_segment<T>(s, s.A + (long long)i * s.ld, margin, c0, c1, lane);
        if (S == 1)
        {
            t = rescue_scan(s, i, t, lane);
            if (lane == 0)
                submit_bid(s, i, t, eps, pmax);
            continue;
        }
        // split row: publish the partial, the last warp to arrive merges all S of them
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
            {
                const Top2<T> o = top2_load_cg(&s.partial[(long long)b * S + k]);
                top2_merge(m, o);
            }
            top2_warp_reduce(m);
            m = rescue_scan(s, i, m, lane);
            if (lane == 0)
            {
                s.arrive[b] = 0;
                submit_bid(s, i, m, eps, pmax);
            }
        }
    }
}

// One thread per bidder.  Exactly one row wins each contested column, so the writes below
// never race; each thread appends at most one row (itself if it lost, else the row it evicted).
template <typename T>
__device__ __forceinline__ void resolve_phase(const DenseState<T>& s, const int* __restrict__ list, int n,
                                              int* __restrict__ next, int* next_count, long long tid,
                                              long long n_threads, int lane)
{
    const long long n_up = ((long long)n + 31) & ~31ll;
    for (long long t = tid; t < n_up; t += n_threads)
    {
        int push = -1;
        if (t < n)
        {
            const int i = list[t];
            const int j = s.bid_col[i];
            if (j >= 0)
            {
                const unsigned long long key = s.slots[j].key;
                if (key == row_key(i))
                {
                    const double bid = s.bid_val[i];
                    s.prices[j] = bid; // Auction.h:259
                    s.pricesf[j] = (float)bid;
                    const int prev = s.owner[j];
                    s.owner[j] = i;
                    s.row2col[i] = j;
                    if (prev >= 0)
                    {
                        s.row2col[prev] = -1; // Auction.h:270
                        push = prev;
                    }
                }
                else
                    push = i;
            }
        }
        const unsigned m = __ballot_sync(kFullMask, push >= 0);
        if (m)
        {
            int base = 0;
            if (lane == 0)
                base = atomicAdd(next_count, __popc(m));
            base = __shfl_sync(kFullMask, base, 0);
            if (push >= 0)
                next[base + __popc(m & ((1u << lane) - 1))] = push;
        }
    }
}

// Start of an eps phase: prices are kept, every assignment is dropped, all rows bid again.
template <typename T>
__device__ __forceinline__ void reset_phase(const DenseState<T>& s, int* __restrict__ list, long long tid,
                                            long long n_threads)
{
    for (long long i = tid; i < s.rows; i += n_threads)
    {
        s.row2col[i] = -1;
        list[i] = (int)i;
    }
    for (long long j = tid; j < s.cols; j += n_threads)
        s.owner[j] = -1;
}

// ---------------------------------------------------------------------------------------------
// Persistent cooperative kernel: the whole solve in one launch, the host is not touched per
// round (north_star (3)).  Launch with cudaLaunchCooperativeKernel, one CTA per SM.
// ---------------------------------------------------------------------------------------------
template <typename T> __global__ void __launch_bounds__(kThreads, 1) k_auction_persistent(DenseState<T> s)
{
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31;
    const int warp_global = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const int total_warps = gridDim.x * kWarpsPerBlock;
    const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
    const long long n_threads = (long long)gridDim.x * kThreads;
    volatile DenseCtrl* ctrl = s.ctrl;

    int cur = 0;
    int phase = 0;
    const int n_phases = ctrl->n_phases;
    const long long max_rounds = ctrl->max_rounds;
    long long rounds = 0, bids = 0;
    double eps = ctrl->eps[0];
    const double amax = ctrl->amax;

    // phase 0 starts from the state the host prepared: list[0] = all rows, count[0] = rows
    for (;;)
    {
        const int n = ctrl->count[cur];
        if (n == 0)
        {
            if (phase + 1 >= n_phases)
                break;
            phase++;
            eps = ctrl->eps[phase];
            grid.sync(); // everybody has read count[cur] == 0 before it is overwritten
            reset_phase(s, s.list[cur], tid, n_threads);
            if (tid == 0)
                ctrl->count[cur] = s.rows;
            grid.sync();
            continue;
        }
        if (rounds >= max_rounds)
        {
            if (tid == 0)
                ctrl->status = 3; // B200LSAP_ERR_MAX_ROUNDS
            break;
        }
        if (tid == 0)
            ctrl->count[cur ^ 1] = 0;
        bid_phase<T>(s, s.list[cur], n, eps, amax, fmax(ctrl->pmax, amax), warp_global, total_warps, lane);
        grid.sync();
        // status is only ever raised inside a bid phase: read here, between the two syncs, every
        // thread sees the same value (reading it at the loop top would race with faster CTAs)
        if (ctrl->status != 0)
            break;
        resolve_phase<T>(s, s.list[cur], n, s.list[cur ^ 1], &s.ctrl->count[cur ^ 1], tid, n_threads, lane);
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
    }
}

// ---------------------------------------------------------------------------------------------
// Host-driven form of the same round (options.driver = 1, and the standalone bidding sweep
// that ncu and bench.py time): identical device code, one launch per phase of a round.
// ---------------------------------------------------------------------------------------------
template <typename T> __global__ void __launch_bounds__(kThreads, 1) k_bid(DenseState<T> s, int cur, int n, double eps)
{
    const double amax = s.ctrl->amax; // written by earlier launches on the stream
    const double pmax = fmax(s.ctrl->pmax, amax);
    const int lane = threadIdx.x & 31;
    const int warp_global = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const int total_warps = gridDim.x * kWarpsPerBlock;
    if (blockIdx.x == 0 && threadIdx.x == 0)
        s.ctrl->count[cur ^ 1] = 0;
    bid_phase<T>(s, s.list[cur], n, eps, amax, pmax, warp_global, total_warps, lane);
}

template <typename T> __global__ void __launch_bounds__(kThreads, 1) k_resolve(DenseState<T> s, int cur, int n)
{
    const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
    const long long n_threads = (long long)gridDim.x * kThreads;
    resolve_phase<T>(s, s.list[cur], n, s.list[cur ^ 1], &s.ctrl->count[cur ^ 1], tid, n_threads, threadIdx.x & 31);
}

template <typename T> __global__ void __launch_bounds__(kThreads, 1) k_reset(DenseState<T> s, int cur)
{
    const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
    const long long n_threads = (long long)gridDim.x * kThreads;
    reset_phase(s, s.list[cur], tid, n_threads);
    if (tid == 0)
        s.ctrl->count[cur] = s.rows;
}

// Initial state of a solve: prices 0 (Auction.h:58), nothing assigned, slots empty.
template <typename T> __global__ void __launch_bounds__(kThreads, 1) k_init(DenseState<T> s, int keep_prices)
{
    const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
    const long long n_threads = (long long)gridDim.x * kThreads;
    reset_phase(s, s.list[0], tid, n_threads);
    for (long long i = tid; i < s.rows; i += n_threads)
    {
        s.arrive[i] = 0;
        s.profits[i] = 1.0; // Auction.h:58
    }
    for (long long j = tid; j < s.ld + 512; j += n_threads)
    {
        const double p = keep_prices && j < s.cols ? s.prices[j] : 0.0;
        s.prices[j] = p;
        s.pricesf[j] = (float)p;
    }
    for (long long j = tid; j < s.cols; j += n_threads)
        s.slots[j] = BidSlot{0ull, 0ull};
    if (tid == 0)
    {
        s.ctrl->count[0] = s.rows;
        s.ctrl->count[1] = 0;
        s.ctrl->status = 0;
        if (!keep_prices)
            s.ctrl->pmax = 0.0;
        s.ctrl->amax = 0.0;
        s.ctrl->phase = 0;
        s.ctrl->rounds = 0;
        s.ctrl->bids = 0;
    }
}

// int32 row2col -> int64 for the host ABI, and price export
__global__ void k_widen_i32(const int* __restrict__ in, long long* __restrict__ out, long long n)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = in[i];
}

} // namespace b200lsap
