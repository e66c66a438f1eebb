// dense_kernels.cuh — the auction for one dense problem on one GPU.
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
// scaled) runs on the device in k_auction_persistent.  Three regimes, chosen by the number of
// bidders n of the round (80 % of the rounds of a square problem have fewer than 8 bidders):
//   n > SMs-1      "wide" round: a warp per row (or per row segment), two cooperative grid syncs
//                  (AuctionMT crosses its ThreadBarrier twice per bid, AuctionMT.h:366-368);
//   n <= SMs-1     "narrow" round: one CTA (or 2, 4) scans one bidder's row, CTA 0 is the
//                  controller: it gathers the n bids, resolves them in shared memory and
//                  publishes the next round; one flag hand-shake per round, no grid sync;
//   n <= tail_rows Gauss-Seidel tail: CTA 0 alone, one bid at a time in the reference's own
//                  schedule, no hand-shake at all.
// n never grows from one round to the next (a bidder wins and evicts at most one row, or loses
// and stays), so a phase moves through the regimes in that order.
#pragma once

#include <cooperative_groups.h>

#include "dense_state.cuh"

namespace b200lsap
{
namespace cg = cooperative_groups;

#ifdef B200LSAP_TIMING
#define TICK(k)                                                                                                    \
    do                                                                                                             \
    {                                                                                                              \
        if (threadIdx.x == 0)                                                                                      \
        {                                                                                                          \
            const long long now_ = clock64();                                                                      \
            sh.dbg[k] += now_ - tick_;                                                                             \
            tick_ = now_;                                                                                          \
        }                                                                                                          \
    } while (0)
#define TICK_INIT long long tick_ = clock64()
#else
#define TICK(k)
#define TICK_INIT
#endif

constexpr int kThreads = 1024;        // one CTA per SM, 32 warps
constexpr int kWarpsPerBlock = kThreads / 32;

// ---------------------------------------------------------------------------------------------
// Warp scan of columns [c0, c1) of one row (c0, c1 multiples of 4, row 16-byte aligned):
// each lane streams 128-bit pieces 512 bytes apart, four in flight.  This is the exact form:
// every element goes through the double-precision top-2 update.  Per-lane result (no reduction).
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
    return t;
}

// ---------------------------------------------------------------------------------------------
// fp32 rows: the same scan behind a single-precision filter.  f = a - (float)p is within
// 2^-23 (amax + pmax) of the exact double a - p, and an element can only change the row's top-2 if
// its exact value exceeds the row's current second best.  `thr` is a warp-uniform lower bound of
// that second best, lowered by `margin` >= 2^-21 (amax + pmax + 1): elements with f <= thr are
// provably irrelevant and cost one FADD and one FMNMX; the few that pass take the exact double
// path of top2_push.  The result is bit-identical to scan_segment_exact; only the instruction
// count and the L1 price traffic (4 B instead of 8 B per element) change.  A lane visits its
// columns in increasing order, as the exact scan does.
//
// The initial threshold comes from the row's candidate list (cand_threshold): any two edges of
// the row bound its second best from below, and the columns that were best at the row's last
// scans are nearly always still the best.
//
// One batch = U 128-bit pieces per lane, piece u at column first + u * stride (a warp scanning a
// row segment: stride 128, first = base + 4 lane; a whole CTA scanning: stride 4 * threads).
// ---------------------------------------------------------------------------------------------
// kTrack: also keep max |a| of everything scanned (the fused sharded solve takes the matrix's amax from its first
// sweep instead of a pass of its own: every row bids in round 1, and with all prices still zero the filter is exact)
template <bool kDeep, bool kTrack = false> struct FilterScan
{
    Top2<float> t;
    float thr;
    float amax;
    __device__ __forceinline__ explicit FilterScan(float thr0) : t(top2_empty<float, kDeep>()), thr(thr0), amax(0.f) {}
};

constexpr int kScanU = 4; // 16-byte loads in flight per lane; 8 is slower (read_probe.cu shows the same for plain reads)

// warp-wide top-2 of per-lane (hi >= lo) pairs of floats; every lane ends with the result
__device__ __forceinline__ void warp_top2f(float& hi, float& lo)
{
#pragma unroll
    for (int sft = 16; sft >= 1; sft >>= 1)
    {
        const float ohi = __shfl_xor_sync(kFullMask, hi, sft);
        const float olo = __shfl_xor_sync(kFullMask, lo, sft);
        lo = fmaxf(fminf(hi, ohi), fmaxf(lo, olo));
        hi = fmaxf(hi, ohi);
    }
}

// the U pieces of one batch, in flight together
__device__ __forceinline__ void scan_load(float4 (&a)[kScanU], const float* __restrict__ row, int first, int stride,
                                          int c1)
{
#pragma unroll
    for (int u = 0; u < kScanU; u++)
    {
        const int cc = first + u * stride;
        a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (cc < c1)
            a[u] = ld_stream4(row + cc);
    }
}

template <bool kDeep, bool kTrack>
__device__ __forceinline__ void scan_consume(FilterScan<kDeep, kTrack>& st, const float4 (&a)[kScanU], const double* prices,
                                             const float* pricesf, float margin, int first, int stride, int c1)
{
    constexpr int U = kScanU;
    float m = -INFINITY;
#pragma unroll
    for (int u = 0; u < U; u++)
    {
        const int cc = first + u * stride;
        if (cc < c1)
        {
            const float4 q = *reinterpret_cast<const float4*>(pricesf + cc);
            m = fmaxf(m, fmaxf(fmaxf(a[u].x - q.x, a[u].y - q.y), fmaxf(a[u].z - q.z, a[u].w - q.w)));
            if (kTrack) // NaNs are ignored, as they are by every comparison of the reference's scan
                st.amax = fmaxf(st.amax, fmaxf(fmaxf(fabsf(a[u].x), fabsf(a[u].y)), fmaxf(fabsf(a[u].z), fabsf(a[u].w))));
        }
    }
    if (!__any_sync(kFullMask, m > st.thr))
        return;
    // Rare path.  First tighten the threshold with the single-precision top-2 of this batch (edges
    // only: a == 0 is "no edge"): the second largest f of the warp, minus the margin, bounds the
    // row's exact second best from below.  Then the survivors take the exact update.
    float hi = -INFINITY, lo = -INFINITY;
#pragma unroll
    for (int u = 0; u < U; u++)
    {
        const int cc = first + u * stride;
        if (cc < c1)
        {
            const float4 q = *reinterpret_cast<const float4*>(pricesf + cc);
            const float f0 = a[u].x != 0.f ? a[u].x - q.x : -INFINITY;
            const float f1 = a[u].y != 0.f ? a[u].y - q.y : -INFINITY;
            const float f2 = a[u].z != 0.f ? a[u].z - q.z : -INFINITY;
            const float f3 = a[u].w != 0.f ? a[u].w - q.w : -INFINITY;
            lo = fmaxf(lo, fminf(hi, f0)), hi = fmaxf(hi, f0);
            lo = fmaxf(lo, fminf(hi, f1)), hi = fmaxf(hi, f1);
            lo = fmaxf(lo, fminf(hi, f2)), hi = fmaxf(hi, f2);
            lo = fmaxf(lo, fminf(hi, f3)), hi = fmaxf(hi, f3);
        }
    }
    warp_top2f(hi, lo);
    const float thr = fmaxf(st.thr, __fsub_rd(lo, margin));
    st.thr = thr;
    // about two elements of the whole warp survive: a warp-uniform branch per piece keeps the
    // sixteen exact updates out of the instruction stream unless some lane needs one
#pragma unroll
    for (int u = 0; u < U; u++)
    {
        const int cc = first + u * stride;
        float4 f = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        if (cc < c1)
        {
            const float4 q = *reinterpret_cast<const float4*>(pricesf + cc);
            f = make_float4(a[u].x - q.x, a[u].y - q.y, a[u].z - q.z, a[u].w - q.w);
        }
        if (__any_sync(kFullMask, fmaxf(fmaxf(f.x, f.y), fmaxf(f.z, f.w)) > thr))
        {
            if (f.x > thr)
                top2_push(st.t, a[u].x, prices[cc], cc);
            if (f.y > thr)
                top2_push(st.t, a[u].y, prices[cc + 1], cc + 1);
            if (f.z > thr)
                top2_push(st.t, a[u].z, prices[cc + 2], cc + 2);
            if (f.w > thr)
                top2_push(st.t, a[u].w, prices[cc + 3], cc + 3);
        }
    }
}

template <bool kDeep, bool kTrack>
__device__ __forceinline__ void scan_batch(FilterScan<kDeep, kTrack>& st, const float* __restrict__ row,
                                           const double* prices, const float* pricesf, float margin, int first,
                                           int stride, int c1)
{
    float4 a[kScanU];
    scan_load(a, row, first, stride, c1);
    scan_consume(st, a, prices, pricesf, margin, first, stride, c1);
}

// Warp-uniform starting threshold of row i from its candidate list (kCand (column, a) pairs left
// by earlier scans of the row; column -1: empty).  Candidates are true edges of the row, so the
// second largest of their current values bounds the row's second best from below whatever their
// age.  fp64 rows are scanned exactly and need none.
template <typename T>
__device__ __forceinline__ float cand_threshold(const DenseState<T>& s, int i, float margin, int lane)
{
    if constexpr (sizeof(T) == 4)
    {
        const int2 raw = __ldcg(s.cand + (long long)i * kCand + lane);
        float hi = raw.x >= 0 ? __int_as_float(raw.y) - s.pricesf[raw.x] : -INFINITY;
        float lo = -INFINITY;
        warp_top2f(hi, lo);
        return __fsub_rd(lo, margin);
    }
    else
        return -INFINITY;
}

// Per-lane scan of columns [c0, c1) of a row by one warp (no reduction).
template <typename T, bool kDeep = false>
__device__ __forceinline__ Top2<T> scan_segment(const DenseState<T>& s, const T* __restrict__ row, float margin,
                                                float thr0, int c0, int c1, int lane)
{
    if constexpr (sizeof(T) == 4)
    {
        FilterScan<kDeep> st(thr0);
        for (int base = c0; base < c1; base += 128 * kScanU) // warp-uniform trip count
            scan_batch(st, row, s.prices, s.pricesf, margin, base + lane * 4, 128, c1);
        return st.t;
    }
    else
        return scan_segment_exact<T, kDeep>(row, s.prices, c0, c1, lane);
}

// The same, also returning max |a| over the lane's elements (fp32 rows only; see FilterScan).
__device__ __forceinline__ Top2<float> scan_segment_amax(const DenseState<float>& s, const float* __restrict__ row,
                                                         float margin, float thr0, int c0, int c1, int lane, float& amax)
{
    FilterScan<false, true> st(thr0);
    for (int base = c0; base < c1; base += 128 * kScanU)
        scan_batch(st, row, s.prices, s.pricesf, margin, base + lane * 4, 128, c1);
    amax = fmaxf(amax, st.amax);
    return st.t;
}

// Per-thread scan of columns [c0, c1) of a row by a whole CTA: thread t takes the 128-bit pieces
// at columns c0 + 4 t + 4 kThreads k.
template <typename T>
__device__ __forceinline__ Top2<T> cta_scan_row(const DenseState<T>& s, const T* __restrict__ row, float margin,
                                                float thr0, int c0, int c1)
{
    if constexpr (sizeof(T) == 4)
    {
        FilterScan<false> st(thr0);
        for (int base = c0; base < c1; base += 4 * kThreads * kScanU)
            scan_batch(st, row, s.prices, s.pricesf, margin, base + (int)threadIdx.x * 4, 4 * kThreads, c1);
        return st.t;
    }
    else
    {
        Top2<T> t = top2_empty<T>();
        for (int c = c0 + (int)threadIdx.x * 4; c < c1; c += 4 * kThreads)
        {
            const auto a = ld_stream4(row + c);
            top2_push(t, a.x, s.prices[c], c);
            top2_push(t, a.y, s.prices[c + 1], c + 1);
            top2_push(t, a.z, s.prices[c + 2], c + 2);
            top2_push(t, a.w, s.prices[c + 3], c + 3);
        }
        return t;
    }
}

// Warp reduction that most warps skip: behind a good threshold only a few lanes of a few warps
// hold a candidate at all.
template <typename T> __device__ __forceinline__ void top2_warp_reduce_sparse(Top2<T>& t)
{
    const unsigned have = __ballot_sync(kFullMask, t.j >= 0);
    if (have == 0)
        return;
    if ((have & (have - 1)) == 0)
    {
        // a single lane holds everything
        const int src = __ffs(have) - 1;
        t.v = __shfl_sync(kFullMask, t.v, src);
        t.w = __shfl_sync(kFullMask, t.w, src);
        t.j = __shfl_sync(kFullMask, t.j, src);
        t.a = __shfl_sync(kFullMask, t.a, src);
        return;
    }
    top2_warp_reduce(t);
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
    top2_warp_reduce(d);
    if (d.j >= 0 && d.w == -CUDART_INF)
        d.w = fmin(-kNoSecondBest, d.v);
    return d;
}

// Lane 0 of the deciding warp: record and offer the bid of row i (wide rounds).
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

template <typename T> __device__ __forceinline__ void cand_store(const DenseState<T>& s, int i, int slot, int j, T a)
{
    if constexpr (sizeof(T) == 4)
        __stcg(s.cand + (long long)i * kCand + slot, make_int2(j, __float_as_int(a)));
}

// ---------------------------------------------------------------------------------------------
// Wide sweeps in half rows.  With more bidders than warps a warp takes whole rows one after the
// other, and the last wave of rows leaves part of the machine idle: 16384 rows on 4736 warps are
// 3.46 rows per warp, the warps with four rows finish alone (a read-only probe of the same access
// pattern loses 11 % to this, tools/micro/read_probe.cu).  Here the two warps 2k, 2k+1 of a CTA
// take the two halves of the same row (6.92 rows per pair: 7 waves, 98.8 % full) and merge their
// top-2 through shared memory: an mbarrier per pair (two arrivals), results double buffered so that
// the next row's result never overwrites one that is still being read.  No global memory, fences or
// atomics, which the general split-row path below needs.
// sweep_split 0: when whole rows would leave the last wave under 95 % full and half rows do
// better; 1: never; 2: whenever there are at least as many bidders as warps.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool pair_rounds(int n, int total_warps, int cols4, int min_seg, int sweep_split)
{
    if (n < total_warps || sweep_split == 1 || cols4 / 2 < min_seg || (total_warps & 1))
        return false;
    if (sweep_split >= 2)
        return true;
    const long long w1 = (n + total_warps - 1) / total_warps, w2 = (2ll * n + total_warps - 1) / total_warps;
    return 100ll * n < 95ll * w1 * total_warps && w2 < 2 * w1;
}

__device__ __forceinline__ unsigned smem_addr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

template <typename T> struct PairShared
{
    unsigned long long bar[kWarpsPerBlock / 2];
    Top2<T> res[kWarpsPerBlock / 2][2];
};

template <typename T>
__device__ __forceinline__ void bid_phase_pairs(const DenseState<T>& s, const int* __restrict__ list, int n, double eps,
                                                float margin, double pmax, int warp_global, int total_warps, int lane)
{
    __shared__ PairShared<T> ps;
    const int wp = (threadIdx.x >> 5) >> 1, half = (threadIdx.x >> 5) & 1;
    __syncthreads();
    if (threadIdx.x < kWarpsPerBlock / 2)
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 2;" ::"r"(smem_addr(&ps.bar[threadIdx.x])));
    __syncthreads();
    int seg = (s.cols4 + 1) / 2;
    seg = (seg + 127) & ~127;
    const int c0 = half * seg, c1 = min(c0 + seg, s.cols4);
    const unsigned bar = smem_addr(&ps.bar[wp]);
    unsigned k = 0;
    for (int b = warp_global >> 1; b < n; b += total_warps >> 1, k++)
    {
        const int i = list[b];
        const float thr0 = cand_threshold(s, i, margin, lane);
        Top2<T> t = scan_segment<T>(s, s.A + (long long)i * s.ld, margin, thr0, c0, c1, lane);
        if (t.j >= 0)
            cand_store(s, i, (t.j >> 2) & (kCand - 1), t.j, t.a);
        top2_warp_reduce_sparse(t);
        if (half == 1 && lane == 0)
            ps.res[wp][k & 1] = t;
        __syncwarp();
        if (lane == 0)
            asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
        unsigned ok;
        do
        {
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                         : "=r"(ok)
                         : "r"(bar), "r"(k & 1)
                         : "memory");
        } while (!ok);
        if (half == 0)
        {
            const Top2<T> o = ps.res[wp][k & 1];
            top2_merge(t, o);
            t = rescue_scan(s, i, t, lane);
            if (lane == 0)
                submit_bid(s, i, t, eps, pmax);
        }
    }
    __syncthreads();
    if (threadIdx.x < kWarpsPerBlock / 2)
        asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_addr(&ps.bar[threadIdx.x])));
}

// (narrow_kernel.cuh, included below)
template <bool kSmem, bool kChase, typename T>
__device__ __forceinline__ bool narrow_eval(const DenseState<T>& s, const double* prices_s, const int* owner_s, int i,
                                            int lane, Top2<T>& t, int& prev, long long* evp);

template <typename T>
__device__ __forceinline__ void bid_phase(const DenseState<T>& s, const int* __restrict__ list, int n, double eps,
                                          double amax, double pmax, int warp_global, int total_warps, int lane)
{
    const float margin = filter_margin(amax, pmax);
    if (s.use_cache && s.ctrl->cache_fresh && !s.dry)
    {
        // Row caches that a narrow phase has just used with hardly a miss (they hold across eps phases: prices only
        // rise) answer the WIDE rounds too: a warp evaluates its bidder's 32 cached columns at the current prices
        // (256 bytes from L2) and, when the second best of them beats the bound on everything else, has the very bid a
        // scan of the 64 KB row would place.  The rare miss scans the row on the spot.  The first sweep of a phase then
        // reads 4 MB instead of the whole matrix.
        for (int b = warp_global; b < n; b += total_warps)
        {
            const int i = list[b];
            Top2<T> t;
            int prev;
            if (!narrow_eval<false, false, T>(s, nullptr, nullptr, i, lane, t, prev, nullptr))
            {
                const float thr0 = cand_threshold(s, i, margin, lane);
                t = scan_segment<T>(s, s.A + (long long)i * s.ld, margin, thr0, 0, s.cols4, lane);
                top2_warp_reduce_sparse(t);
                t = rescue_scan(s, i, t, lane);
            }
            if (lane == 0)
                submit_bid(s, i, t, eps, pmax);
        }
        return;
    }
    if (pair_rounds(n, total_warps, s.cols4, s.min_seg, s.sweep_split))
    {
        bid_phase_pairs<T>(s, list, n, eps, margin, pmax, warp_global, total_warps, lane);
        return;
    }
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
        // Without a starting threshold a scan takes the exact path in its first batch and again at
        // every batch that beats its running second best (about 2 ln(batches) times: 7 of a whole
        // row's 32 batches); with the candidates' threshold only the batches that hold the row's
        // best columns do.  Measured on 16384 x 16384: 5.2 -> 6.0 TB/s for whole rows.
        const float thr0 = cand_threshold(s, i, margin, lane);
        Top2<T> t = scan_segment<T>(s, s.A + (long long)i * s.ld, margin, thr0, c0, c1, lane);
        if (S == 1)
        {
            if (t.j >= 0)
                cand_store(s, i, (t.j >> 2) & (kCand - 1), t.j, t.a); // = lane: next scan's candidates
            top2_warp_reduce_sparse(t);
            t = rescue_scan(s, i, t, lane);
            if (lane == 0)
                submit_bid(s, i, t, eps, pmax);
            continue;
        }
        // split row: publish the partial, the last warp to arrive merges all S of them
        top2_warp_reduce_sparse(t);
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

// ---------------------------------------------------------------------------------------------
// Narrow rounds and the Gauss-Seidel tail (n <= SMs - 1 bidders).
//
// CTA 0 is the controller, CTAs 1.. are scanners.  Bidder b's row is scanned by P = 1, 2 or 4
// whole CTAs (P column parts; scanner w serves bidder w / P, part w % P), which merge their 32
// warps through shared memory and publish one record.  The controller waits for the n * P
// records, merges the parts (the thread-0 merge of AuctionMT.h:377-398), resolves the n bids in
// shared memory (highest bid per column, lowest row among equals: the rule of the wide rounds),
// applies prices / owners / evictions, and posts the next round into the scanners' mailboxes.
// Rounds with at most tail_rows bidders are finished by the controller alone, Gauss-Seidel.
// The list is put in increasing row order at entry, so the outcome does not depend on the
// order in which the wide rounds' atomics compacted it.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int ld_volatile(const int* p) { return *reinterpret_cast<const volatile int*>(p); }
__device__ __forceinline__ unsigned ld_volatile(const unsigned* p)
{
    return *reinterpret_cast<const volatile unsigned*>(p);
}

__device__ __forceinline__ void prefetch_row_l2(const void* row, unsigned bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(row), "r"(bytes) : "memory");
}

template <typename T> struct NarrowShared
{
    Top2<T> part[kWarpsPerBlock];
    int q[kNarrowMax];      // controller: bidders of the round, then the next list
    int col[kNarrowMax];    // controller: column each bidder bid for
    double bid[kNarrowMax]; // controller: its bid
    int warp_count[kWarpsPerBlock];
    int head, qn, stop;
    double pmax;
    long long dbg[16]; // cycle counters (B200LSAP_TIMING builds)
};

// Every thread of a CTA: scan columns [c0, c1) of row i, merge the CTA's warps.  The result is
// valid in all lanes of warp 0.  Also refreshes the row's candidate list.
template <typename T>
__device__ __forceinline__ Top2<T> cta_bid_scan(const DenseState<T>& s, NarrowShared<T>& sh, int i, float margin, int c0,
                                                int c1)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#ifdef B200LSAP_TIMING
    long long tick_ = clock64();
    const bool timed_ = blockIdx.x == 0;
#define TICKB(k)                                                                                                   \
    do                                                                                                             \
    {                                                                                                              \
        if (timed_ && threadIdx.x == 0)                                                                            \
        {                                                                                                          \
            const long long now_ = clock64();                                                                      \
            sh.dbg[k] += now_ - tick_;                                                                             \
            tick_ = now_;                                                                                          \
        }                                                                                                          \
    } while (0)
#else
#define TICKB(k)
#endif
    const float thr0 = cand_threshold(s, i, margin, lane);
    TICKB(8);
    Top2<T> t = cta_scan_row(s, s.A + (long long)i * s.ld, margin, thr0, c0, c1);
    TICKB(9);
    top2_warp_reduce_sparse(t);
    if (lane == 0)
        sh.part[warp] = t;
    TICKB(10);
    __syncthreads();
    TICKB(11);
    Top2<T> m = top2_empty<T>();
    if (warp == 0)
    {
        m = sh.part[lane];
        // candidates: a column always lives in slot (column / 4) % kCand, so the list never holds
        // an edge twice; of the warp results that share a slot the best one is kept
        if constexpr (sizeof(T) == 4)
        {
            const int slot = m.j >= 0 ? (m.j >> 2) & (kCand - 1) : -1 - lane;
            bool keep = m.j >= 0;
            for (int k = 0; k < 32; k++)
            {
                const double ov = __shfl_sync(kFullMask, m.v, k);
                const int oslot = __shfl_sync(kFullMask, slot, k);
                keep = keep && !(oslot == slot && (ov > m.v || (ov == m.v && k < lane)));
            }
            if (keep)
                cand_store(s, i, slot, m.j, m.a);
        }
        top2_warp_reduce_sparse(m);
    }
    return m;
}

// Scanner CTAs.  `last` is the last mailbox flag this CTA has consumed (kept across phases).
template <typename T>
__device__ __forceinline__ void narrow_scanner(const DenseState<T>& s, NarrowShared<T>& sh, int& last, double amax)
{
    NarrowSlot* slot = &s.narrow->slot[blockIdx.x - 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (;;)
    {
        if (threadIdx.x == 0)
        {
            int f;
            while ((f = ld_volatile(&slot->flag)) == last)
                __nanosleep(32);
            __threadfence();
            sh.head = f;
        }
        __syncthreads();
        last = sh.head;
        if (last < 0)
            return; // the controller finished the phase (or stopped on an error)
        const int i = ld_volatile(&slot->row);
        const int c0 = ld_volatile(&slot->c0), c1 = ld_volatile(&slot->c1);
        const float margin = filter_margin(amax, *reinterpret_cast<const volatile double*>(&slot->pmax));
        Top2<T> m = cta_bid_scan(s, sh, i, margin, c0, c1);
        if (warp == 0 && lane == 0)
        {
            slot->v = m.v;
            slot->w = m.w;
            slot->a = (double)m.a;
            slot->j = m.j;
            __threadfence();
            atomicAdd(&s.narrow->arrived, 1u);
        }
        __syncthreads(); // sh.part / sh.head are reused next round
    }
}

// Gauss-Seidel tail on the controller: the queue sh.q[head .. head+qn) is served FIFO, one bid at a
// time (Auction.h:220-285), until it is empty.  Returns the bids placed.
template <typename T>
__device__ __forceinline__ long long gs_tail(const DenseState<T>& s, NarrowShared<T>& sh, double eps, double amax,
                                             long long budget)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned row_bytes = (unsigned)(s.cols4 * sizeof(T));
    const bool can_prefetch = (row_bytes & 15u) == 0 && ((s.ld * sizeof(T)) & 15u) == 0;
    if (tid > 0 && tid < sh.qn && can_prefetch)
        prefetch_row_l2(s.A + (long long)sh.q[(sh.head + tid) & (kNarrowMax - 1)] * s.ld, row_bytes);
    long long done = 0;
    for (;;)
    {
        const int qn = sh.qn, head = sh.head;
        if (qn == 0 || sh.stop || done >= budget)
            break;
        const int i = sh.q[head & (kNarrowMax - 1)];
        const float margin = filter_margin(amax, sh.pmax);
        TICK_INIT;
        Top2<T> m = cta_bid_scan(s, sh, i, margin, 0, s.cols4);
        TICK(0);
        if (warp == 0)
        {
            m = rescue_scan(s, i, m, lane);
            if (lane == 0)
            {
                if (m.j < 0)
                {
                    s.ctrl->status = 2; // B200LSAP_ERR_INFEASIBLE
                    sh.stop = 1;
                }
                else
                {
                    const double bid = (double)m.a - m.w + eps; // Auction.h:250
                    s.profits[i] = m.w - eps;                   // Auction.h:253
                    s.prices[m.j] = bid;                        // Auction.h:259
                    s.pricesf[m.j] = (float)bid;
                    const int prev = s.owner[m.j];
                    s.owner[m.j] = i;
                    s.row2col[i] = m.j;
                    int left = qn - 1;
                    if (prev >= 0)
                    {
                        s.row2col[prev] = -1; // Auction.h:270
                        sh.q[(head + qn) & (kNarrowMax - 1)] = prev;
                        left++;
                        if (can_prefetch && left > 1)
                            prefetch_row_l2(s.A + (long long)prev * s.ld, row_bytes);
                    }
                    sh.head = head + 1;
                    sh.qn = left;
                    if (bid > sh.pmax)
                        sh.pmax = bid;
                }
            }
        }
        TICK(1);
        __syncthreads();
        TICK(2);
        done++;
    }
    return done;
}

struct NarrowCounts
{
    long long rounds, bids, gs_bids, gs_runs;
};

// Controller CTA: finishes the eps phase that list[0..n) is the tail of.  On return the list is
// empty unless the watchdog or an error stopped it (then ctrl->status is set).
template <typename T>
__device__ __forceinline__ void narrow_controller(const DenseState<T>& s, NarrowShared<T>& sh,
                                                  const int* __restrict__ list, int n, int& seq, int mode_count,
                                                  double eps, double amax, long long budget, NarrowCounts& cnt)
{
    NarrowCtrl* nc = s.narrow;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int scanners = gridDim.x - 1;
    // entry: bidders in increasing row order (rank sort, n <= kNarrowMax)
    if (tid < n)
    {
        const int mine = list[tid];
        int rank = 0;
        for (int k = 0; k < n; k++)
            rank += list[k] < mine;
        sh.q[rank] = mine;
    }
    if (tid == 0)
    {
        sh.head = 0;
        sh.qn = n;
        sh.stop = 0;
        sh.pmax = fmax(*reinterpret_cast<volatile double*>(&s.ctrl->pmax), amax);
    }
    if (tid < 16)
        sh.dbg[tid] = 0;
    __syncthreads();
    unsigned target = ld_volatile(&nc->arrived);
    const long long spent0 = cnt.rounds + cnt.gs_bids;
    for (;;)
    {
        n = sh.qn;
        if (n == 0 || sh.stop)
            break;
        const long long left_budget = budget - (cnt.rounds + cnt.gs_bids - spent0);
        if (left_budget <= 0)
        {
            if (tid == 0)
                s.ctrl->status = 3; // B200LSAP_ERR_MAX_ROUNDS
            break;
        }
        if (n <= s.tail_rows || n > s.narrow_rows || n > scanners)
        {
            cnt.gs_bids += gs_tail(s, sh, eps, amax, left_budget);
            cnt.gs_runs++;
            if (sh.qn != 0 && !sh.stop && tid == 0)
                s.ctrl->status = 3; // stopped on the watchdog
            break;
        }
        // ---- publish the round (the queue is compacted to q[0..n) at this point, head == 0) -----
        int parts = 1;
        while (parts < s.max_parts && n * parts * 2 <= scanners && (s.cols4 + parts * 2 - 1) / (parts * 2) >= 1024)
            parts *= 2;
        TICK_INIT;
        ++seq;
        if (tid < n * parts)
        {
            const int b = tid / parts, part = tid - b * parts;
            int seg = (s.cols4 + parts - 1) / parts;
            seg = (seg + 127) & ~127;
            NarrowSlot* slot = &nc->slot[tid];
            slot->row = sh.q[b];
            slot->c0 = part * seg;
            slot->c1 = min(part * seg + seg, s.cols4);
            slot->pmax = sh.pmax;
            __threadfence();
            *reinterpret_cast<volatile int*>(&slot->flag) = seq;
        }
        if (tid == 0)
        {
            target += (unsigned)(n * parts);
            while ((int)(ld_volatile(&nc->arrived) - target) < 0)
                ;
            __threadfence();
        }
        TICK(4);
        __syncthreads();
        // ---- gather: one thread per bidder merges its parts (AuctionMT.h:377-398) ---------------
        int my_row = -1, my_col = -1, my_prev = -1;
        double my_bid = 0.0, my_w = 0.0;
        if (tid < n)
        {
            my_row = sh.q[tid];
            Top2<double> m;
            m.v = m.w = -kNoSecondBest;
            m.j = -1;
            m.a = 0.0;
            for (int p = 0; p < parts; p++)
            {
                const NarrowSlot* r = &nc->slot[tid * parts + p];
                Top2<double> o;
                o.v = __ldcg(&r->v);
                o.w = __ldcg(&r->w);
                o.a = __ldcg(&r->a);
                o.j = __ldcg(&r->j);
                top2_merge(m, o);
            }
            if (m.j < 0)
                sh.stop = 2; // nothing above the -1e6 sentinel: needs the deep re-scan of the tail
            else
            {
                my_col = m.j;
                my_w = m.w;
                my_bid = m.a - m.w + eps; // Auction.h:250; m.a is the exactly widened coefficient
            }
            sh.col[tid] = my_col;
            sh.bid[tid] = my_bid;
        }
        __syncthreads();
        TICK(5);
        if (sh.stop == 2)
        {
            // hand the whole queue to the Gauss-Seidel tail, which knows the rescue path
            __syncthreads();
            if (tid == 0)
                sh.stop = 0;
            __syncthreads();
            cnt.gs_bids += gs_tail(s, sh, eps, amax, left_budget);
            cnt.gs_runs++;
            if (sh.qn != 0 && !sh.stop && tid == 0)
                s.ctrl->status = 3;
            break;
        }
        // ---- resolve in shared memory: highest bid per column, lowest row among equal bids -------
        bool win = false;
        if (tid < n)
        {
            win = true;
            for (int k = 0; k < n; k++)
                if (k != tid && sh.col[k] == my_col)
                {
                    const double ob = sh.bid[k];
                    const int orow = sh.q[k];
                    if (ob > my_bid || (ob == my_bid && orow < my_row))
                        win = false;
                }
            s.profits[my_row] = my_w - eps; // Auction.h:253
            if (win)
            {
                s.prices[my_col] = my_bid; // Auction.h:259
                s.pricesf[my_col] = (float)my_bid;
                my_prev = s.owner[my_col];
                s.owner[my_col] = my_row;
                s.row2col[my_row] = my_col;
                if (my_prev >= 0)
                    s.row2col[my_prev] = -1; // Auction.h:270
            }
        }
        // ---- next list, in bidder order: a loser stays, a winner hands its place to the row it evicted
        const int push = tid < n ? (win ? my_prev : my_row) : -1;
        const unsigned mask = __ballot_sync(kFullMask, push >= 0);
        if (lane == 0)
            sh.warp_count[warp] = __popc(mask);
        double bmax = (tid < n && win) ? my_bid : 0.0;
#pragma unroll
        for (int sft = 16; sft >= 1; sft >>= 1)
            bmax = fmax(bmax, __shfl_xor_sync(kFullMask, bmax, sft));
        __syncthreads(); // everybody has read sh.q / sh.col / sh.bid, warp_count is complete
        int base = 0, total = 0;
        for (int k = 0; k < kWarpsPerBlock; k++)
        {
            const int c = sh.warp_count[k];
            base += k < warp ? c : 0;
            total += c;
        }
        if (push >= 0)
            sh.q[base + __popc(mask & ((1u << lane) - 1))] = push;
        if (lane == 0 && bmax > 0.0)
            atomicMax(reinterpret_cast<unsigned long long*>(&sh.pmax), (unsigned long long)__double_as_longlong(bmax));
        if (tid == 0)
        {
            sh.head = 0;
            sh.qn = total;
        }
        cnt.rounds++;
        cnt.bids += n;
        __syncthreads();
        TICK(6);
    }
    // release the scanners: a negative flag, different from the one of the previous narrow mode
    __syncthreads();
    if (tid < scanners)
    {
        __threadfence();
        *reinterpret_cast<volatile int*>(&nc->slot[tid].flag) = -mode_count;
    }
    if (tid == 0 && sh.pmax > s.ctrl->pmax)
        s.ctrl->pmax = sh.pmax;
#ifdef B200LSAP_TIMING
    if (tid < 16)
        s.ctrl->dbg[tid] += sh.dbg[tid];
#endif
}

} // namespace b200lsap
#include "dense_cache.cuh"
#include "narrow_kernel.cuh"
namespace b200lsap
{

// One CTA plays alone for a while (a narrow episode) and the others have nothing to do until it is done.  Parked at a
// grid sync they poll the barrier word back to back — one thread per CTA, 147 of them on ONE L2 address, with an L1
// invalidate per poll (ncu on the sparse solve: 95 % of all warp samples in that spin).  Here they doze instead, polling
// once a microsecond, and go to the grid sync when the episode is over.  Measured on the sparse 20000 x 20000 case: no
// change in solve time (163 ms either way: the working CTA is bound by its own chain of L2 round trips, not by the
// pollers), so this only saves the power and L2 traffic of 147 spinning SMs.
__device__ __forceinline__ void solo_wait(DenseCtrl* ctrl, int seq)
{
    if (threadIdx.x == 0)
        while (*reinterpret_cast<volatile int*>(&ctrl->solo_seq) < seq)
            __nanosleep(1000);
    __syncthreads();
}
__device__ __forceinline__ void solo_done(DenseCtrl* ctrl, int seq)
{
    __syncthreads();
    if (threadIdx.x == 0)
    {
        __threadfence();
        *reinterpret_cast<volatile int*>(&ctrl->solo_seq) = seq;
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
template <typename T>
__global__ void __launch_bounds__(kThreads, 1) k_auction_persistent(const __grid_constant__ DenseState<T> s)
{
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31;
    const int warp_global = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const int total_warps = gridDim.x * kWarpsPerBlock;
    const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
    const long long n_threads = (long long)gridDim.x * kThreads;
    volatile DenseCtrl* ctrl = s.ctrl;
    __shared__ NarrowShared<T> narrow_sh;
    __shared__ BuildShared<T> build_sh[SweepGroups<T>::count];

    // The fp32 solve runs as one launch per eps phase (k_narrow finishes each phase's narrow rounds in
    // between, narrow_kernel.cuh): the loop state lives in the controller block across launches.
    if (ctrl->done || ctrl->status != 0)
        return;
#ifdef B200LSAP_TIMING
    const long long tk0 = clock64();
#endif
    int cur = ctrl->cur;
    int phase = ctrl->phase;
    int seq = 0;        // CTA 0: sequence number of the last narrow round; scanners: last mailbox flag consumed
    int mode_count = 0; // times the grid has entered narrow mode
    const int n_phases = ctrl->n_phases;
    const long long max_rounds = ctrl->max_rounds - ctrl->rounds;
    long long rounds = 0, bids = 0;
    NarrowCounts ncnt{0, 0, 0, 0};
    double eps = ctrl->eps[phase < 0 ? 0 : phase]; // phase == -1: k_narrow restarted the schedule (escalate_schedule)
    const double amax = ctrl->amax;
    long long phase_rounds = 0; // wide rounds of the current phase in this launch
    int small_rounds = 0;       // of those, rounds that could have been narrow (see narrow_after)
    const bool cached = s.use_cache != 0;
    const int narrow_rows = cached ? min(s.narrow_rows, kNarrowMax) : min(min(s.narrow_rows, (int)gridDim.x - 1), kNarrowMax);
    bool handed_over = false;

    // phase 0 starts from the state the host prepared: list[0] = all rows, count[0] = rows
    for (;;)
    {
        const int n = ctrl->count[cur];
        if (n == 0 || (n <= s.phase_cut && phase >= 0 && phase + 1 < n_phases))
        {
            if (n == 0 && s.asym_repair && phase >= 0 && phase + s.asym_repair >= n_phases)
            {
                // End of a phase on rows < cols: no unassigned object may be priced above an assigned one
                // (AuctionCommon.h:347-354 with lambda = the lowest assigned price; what the reference's reverse auction
                // is for, Auction.h:346-469).  When one is, every price below the highest unassigned price F rises to F —
                // the collective move of the cols - rows dummy bidders of the equivalent square problem; prices stay
                // monotone, so candidate lists, row caches and bid slots stay valid — and the rows whose object was
                // raised bid again AT THIS PHASE'S eps.  Done per phase, the last phase starts from prices whose floor
                // is already right and has little left to repair.
                if (tid == 0)
                {
                    ctrl->lam_hi = 0ull;
                    ctrl->lam_lo = (unsigned long long)__double_as_longlong(1e300);
                }
                grid.sync();
                double hi = 0.0, lo = 1e300;
                for (long long j = tid; j < s.cols; j += n_threads)
                {
                    const double p = s.prices[j];
                    if (s.owner[j] < 0)
                        hi = fmax(hi, p);
                    else
                        lo = fmin(lo, p);
                }
#pragma unroll
                for (int sft = 16; sft >= 1; sft >>= 1)
                {
                    hi = fmax(hi, __shfl_xor_sync(kFullMask, hi, sft));
                    lo = fmin(lo, __shfl_xor_sync(kFullMask, lo, sft));
                }
                if (lane == 0)
                {
                    // non-negative doubles order like their bits
                    atomicMax(&s.ctrl->lam_hi, (unsigned long long)__double_as_longlong(hi));
                    atomicMin(&s.ctrl->lam_lo, (unsigned long long)__double_as_longlong(lo));
                }
                grid.sync();
                const double F = __longlong_as_double((long long)ctrl->lam_hi), L = __longlong_as_double((long long)ctrl->lam_lo);
                if (F > L)
                {
                    for (long long j = tid; j < s.cols; j += n_threads)
                        if (s.prices[j] < F)
                        {
                            s.prices[j] = F;
                            s.pricesf[j] = (float)F;
                            const int o = s.owner[j];
                            if (o >= 0)
                            {
                                s.row2col[o] = -1;
                                s.owner[j] = -1;
                            }
                        }
                    if (tid == 0)
                        ctrl->repairs = ctrl->repairs + 1;
                    grid.sync();
                    const long long up = ((long long)s.rows + 31) & ~31ll;
                    for (long long i = tid; i < up; i += n_threads)
                    {
                        const bool push = i < s.rows && s.row2col[i] < 0;
                        const unsigned m = __ballot_sync(kFullMask, push);
                        if (m)
                        {
                            int base = 0;
                            if (lane == 0)
                                base = atomicAdd(&s.ctrl->count[cur], __popc(m));
                            base = __shfl_sync(kFullMask, base, 0);
                            if (push)
                                s.list[cur][base + __popc(m & ((1u << lane) - 1))] = (int)i;
                        }
                    }
                    grid.sync();
                    small_rounds = 0;
                    continue; // the same phase goes on with the unassigned rows
                }
            }
            if (phase + 1 >= n_phases)
                break;
            phase++;
            eps = ctrl->eps[phase];
            phase_rounds = 0;
            small_rounds = 0;
            grid.sync(); // everybody has read count[cur] == 0 before it is overwritten
            reset_phase(s, s.list[cur], tid, n_threads);
            if (tid == 0)
            {
                ctrl->count[cur] = s.rows;
                ctrl->phase_work = 0;
            }
            grid.sync();
            continue;
        }
        const long long spent = rounds + ncnt.rounds + ncnt.gs_bids;
        if (spent >= max_rounds)
        {
            if (tid == 0)
                ctrl->status = 3; // B200LSAP_ERR_MAX_ROUNDS
            break;
        }
        bool go_narrow = n <= narrow_rows || n <= min(s.tail_rows, kNarrowMax);
        if (go_narrow && cached && small_rounds < s.narrow_after)
        {
            go_narrow = false; // not yet worth a cache-building sweep of the whole matrix (DenseState::narrow_after)
            small_rounds++;
        }
        if (go_narrow)
        {
            // narrow rounds / Gauss-Seidel tail: finishes the phase without a grid sync per round
            mode_count++;
            if (cached && n <= narrow_rows)
            {
                // a fresh cache for every row, then this launch ends: k_narrow (one SM, prices in
                // shared memory) finishes the phase, the next launch starts the next one
#ifdef B200LSAP_TIMING
                const long long tb0 = clock64();
#endif
                // (the sweep reads the whole matrix: 0.36 ms at 16384 x 16384, once per phase.  Caches stay VALID across
                // phases — prices only rise, so every bound holds — and in the later phases eps is far below the spread
                // of a row's 32 best values, so they also stay GOOD: the sweep is skipped while the last narrow phase
                // hardly missed, k_narrow re-scans the few rows that do miss)
                if (!ctrl->cache_fresh || s.rebuild_caches_every_phase)
                    cache_build_sweep<T>(s, build_sh, filter_margin(amax, fmax(ctrl->pmax, amax)));
#ifdef B200LSAP_TIMING
                if (tid == 0)
                    s.ctrl->dbg[14] += clock64() - tb0;
#endif
                handed_over = true;
                break;
            }
            if (blockIdx.x == 0)
                narrow_controller<T>(s, narrow_sh, s.list[cur], n, seq, mode_count, eps, amax, max_rounds - spent, ncnt);
            else
                narrow_scanner<T>(s, narrow_sh, seq, amax);
            grid.sync();
            if (ctrl->status != 0)
                break;
            if (tid == 0)
                ctrl->count[cur] = 0;
            grid.sync();
            continue;
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
        phase_rounds++;
        bids += n;
    }
    if (tid == 0)
    {
        ctrl->phase_work += phase_rounds;
        ctrl->rounds += rounds + ncnt.rounds;
        ctrl->bids += bids + ncnt.bids + ncnt.gs_bids;
        ctrl->narrow_rounds += ncnt.rounds;
        ctrl->gs_bids += ncnt.gs_bids;
        ctrl->gs_episodes += ncnt.gs_runs;
        ctrl->phase = phase;
        ctrl->cur = cur;
        if (!handed_over)
            ctrl->done = 1;
#ifdef B200LSAP_TIMING
        s.ctrl->dbg[15] += clock64() - tk0;
        s.ctrl->dbg[13] += rounds;
#endif
    }
}

// ---------------------------------------------------------------------------------------------
// Host-driven form of the wide round (options.driver = 1, and the standalone bidding sweep
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
    if (keep_prices < 2) // 2: same matrix as before; candidates are edges of their rows whatever the prices
        for (long long k = tid; k < (long long)s.rows * kCand; k += n_threads)
            s.cand[k] = make_int2(-1, 0);
    if (s.cache_hdr)
    {
        for (long long i = tid; i < s.rows; i += n_threads)
            s.cache_hdr[i] = RowCacheHdr{0.0, 0, 0};
        for (long long k = tid; k < (long long)s.rows * kCacheK; k += n_threads)
            s.cache[k] = make_ent<T>(0, T(0)); // every slot always names a real column (narrow_eval)
    }
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
        s.ctrl->cache_hits = s.ctrl->cache_misses = s.ctrl->cache_fallbacks = 0;
        s.ctrl->narrow_rounds = s.ctrl->gs_bids = s.ctrl->gs_episodes = 0;
        s.ctrl->cur = 0;
        s.ctrl->done = 0;
        s.ctrl->phase_work = 0;
        s.ctrl->escalations = 0;
        s.ctrl->esc_phase = 0;
        s.narrow->arrived = 0u;
    }
    if (tid < kNarrowMax)
        s.narrow->slot[tid].flag = 0;
}

// int32 row2col -> int64 for the host ABI, and price export
__global__ void k_widen_i32(const int* __restrict__ in, long long* __restrict__ out, long long n)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = in[i];
}

} // namespace b200lsap
