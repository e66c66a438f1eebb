// dense_cache.cuh — row caches: narrow rounds without row scans (fp32 problems).
//
// Included by dense_kernels.cuh (needs NarrowShared, cta_bid_scan, rescue_scan).
//
// A square problem spends nearly all of its ROUNDS with a handful of bidders, and the same rows bid
// again and again against prices that moved a little.  findForward (Auction.h:124-152) re-reads the
// whole row for every bid; what it returns depends only on the few columns near the top of the row.
//
//   cache[row]      the kCacheK best (column, a) pairs of the row at its last CTA scan, best first
//   hdr.bound       an upper bound of a - p over every column that is NOT cached, at that time
//
// Prices never decrease (a price only ever becomes a winning bid, which exceeds it), so a - p of a
// column never increases and `bound` stays an upper bound for ever.  A bidder evaluates its cached
// columns exactly, (double)a - p with the current prices and the reference's tie rules; when the
// second best of them exceeds `bound`, no other column of the row can be best or second best and
// the bid a - w + eps (Auction.h:250) is bit-identical to the one a full scan would place.
// Otherwise ("miss") the row is scanned again and its cache rebuilt.
//
// The CTA scan that builds a cache (cta_scan_build) is a threshold selection, not a running top-2:
//   1. every thread loads its pieces of the row and f = a - (float)p; the CTA takes the top-2 of
//      each warp's elements (64 values);
//   2. T = the 32nd largest of those 64 (each is a distinct element, so >= 32 elements reach it),
//      lowered by twice the filter margin;
//   3. elements with f > T ("survivors", ~36 on uniform data) are compacted into shared memory;
//   4. survivors are evaluated exactly in double and ranked (value descending, column ascending:
//      strict '>' in column order, Auction.h:136); rank 0 / 1 are the row's best / second best,
//      ranks < kCacheK go to the cache, bound = max(T + margin, value of rank kCacheK).
// An element that did not survive has f <= T, hence exact value < T + margin (the filter's error
// bound, dense_state.cuh filter_margin).  The scan answers only when its exact second best exceeds
// T + margin; otherwise (dense ties, fewer than two edges, more than kSurvCap survivors) it reports
// failure, the row keeps no cache and the caller uses the unfiltered path (cta_bid_scan + rescue).
#pragma once

#ifndef B200LSAP_VAR
#define B200LSAP_VAR 0
#endif

namespace b200lsap
{

struct BuildShared
{
    float cand[64];        // warp w: cand[w] = largest f of its elements, cand[32 + w] = second largest
    int2 surv[kSurvCap];   // survivors: (column, bits of a)
    double sval[kSurvCap]; // their exact values
    int cnt;
    int ok;
    float T;
    int j;
    float a;
    double v, w, kth;
};

constexpr int kBuildRegPieces = 4; // pieces of a row a thread keeps in registers between the two passes

__device__ __forceinline__ float4 piece_f(const float4 a, const float4 q)
{
    return make_float4(a.x != 0.f ? a.x - q.x : -INFINITY, a.y != 0.f ? a.y - q.y : -INFINITY,
                       a.z != 0.f ? a.z - q.z : -INFINITY, a.w != 0.f ? a.w - q.w : -INFINITY);
}

// Every thread of the CTA.  Scans row i, rebuilds its cache, returns true when the result
// (bs.v, bs.w, bs.j, bs.a) is exact; all threads see the same return value and may read bs until
// the next call.  On false the row's cache is left empty.
__device__ __forceinline__ bool cta_scan_build(const DenseState<float>& s, BuildShared& bs, int i, float margin)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* __restrict__ row = s.A + (long long)i * s.ld;
    const int c1 = s.cols4;
    constexpr int stride = 4 * kThreads;
    float4 areg[kBuildRegPieces];
    // ---- pass 1: f = a - (float)p, top-2 of the thread's elements -------------------------------
    float hi = -INFINITY, lo = -INFINITY;
#pragma unroll
    for (int u = 0; u < kBuildRegPieces; u++)
    {
        const int cc = tid * 4 + u * stride;
        areg[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (cc < c1)
            areg[u] = ld_stream4(row + cc);
    }
#pragma unroll
    for (int u = 0; u < kBuildRegPieces; u++)
    {
        const int cc = tid * 4 + u * stride;
        if (cc < c1)
        {
            const float4 f = piece_f(areg[u], *reinterpret_cast<const float4*>(s.pricesf + cc));
            lo = fmaxf(lo, fminf(hi, f.x)), hi = fmaxf(hi, f.x);
            lo = fmaxf(lo, fminf(hi, f.y)), hi = fmaxf(hi, f.y);
            lo = fmaxf(lo, fminf(hi, f.z)), hi = fmaxf(hi, f.z);
            lo = fmaxf(lo, fminf(hi, f.w)), hi = fmaxf(hi, f.w);
        }
    }
    for (int cc = tid * 4 + kBuildRegPieces * stride; cc < c1; cc += stride) // rows longer than 16384 columns
    {
        const float4 f = piece_f(ld_stream4(row + cc), *reinterpret_cast<const float4*>(s.pricesf + cc));
        lo = fmaxf(lo, fminf(hi, f.x)), hi = fmaxf(hi, f.x);
        lo = fmaxf(lo, fminf(hi, f.y)), hi = fmaxf(hi, f.y);
        lo = fmaxf(lo, fminf(hi, f.z)), hi = fmaxf(hi, f.z);
        lo = fmaxf(lo, fminf(hi, f.w)), hi = fmaxf(hi, f.w);
    }
    warp_top2f(hi, lo);
    if (lane == 0)
    {
        bs.cand[warp] = hi;
        bs.cand[32 + warp] = lo;
    }
    if (tid == 0)
    {
        bs.cnt = 0;
        bs.ok = 0;
        bs.T = INFINITY;
        bs.kth = -CUDART_INF;
    }
    __syncthreads();
    // ---- threshold: the r-th largest of the (up to) 64 candidates, r = 32 for a full CTA --------
    if (tid < 64)
    {
        const float x = bs.cand[tid];
        int rank = 0, finite = 0;
        for (int u = 0; u < 64; u++)
        {
            const float y = bs.cand[u];
            rank += (y > x) || (y == x && u < tid);
            finite += y > -INFINITY;
        }
        const int r = min(32, max(2, (finite + 1) / 2));
        if (finite >= 2 && rank == r - 1)
            bs.T = x - 2.f * margin;
    }
    __syncthreads();
    const float T = bs.T;
    if (!(T < INFINITY)) // fewer than two edges in the row
    {
        if (tid == 0)
            s.cache_hdr[i] = RowCacheHdr{0.0, 0, 0};
        __syncthreads();
        return false;
    }
    // ---- pass 2: compact the survivors ----------------------------------------------------------
#pragma unroll
    for (int u = 0; u < kBuildRegPieces; u++)
    {
        const int cc = tid * 4 + u * stride;
        if (cc < c1)
        {
            const float4 f = piece_f(areg[u], *reinterpret_cast<const float4*>(s.pricesf + cc));
            if (fmaxf(fmaxf(f.x, f.y), fmaxf(f.z, f.w)) > T)
            {
                const float fv[4] = {f.x, f.y, f.z, f.w};
                const float av[4] = {areg[u].x, areg[u].y, areg[u].z, areg[u].w};
#pragma unroll
                for (int e = 0; e < 4; e++)
                    if (fv[e] > T)
                    {
                        const int slot = atomicAdd(&bs.cnt, 1);
                        if (slot < kSurvCap)
                            bs.surv[slot] = make_int2(cc + e, __float_as_int(av[e]));
                    }
            }
        }
    }
    for (int cc = tid * 4 + kBuildRegPieces * stride; cc < c1; cc += stride) // re-read from L2
    {
        const float4 a4 = ld_stream4(row + cc);
        const float4 f = piece_f(a4, *reinterpret_cast<const float4*>(s.pricesf + cc));
        if (fmaxf(fmaxf(f.x, f.y), fmaxf(f.z, f.w)) > T)
        {
            const float fv[4] = {f.x, f.y, f.z, f.w};
            const float av[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
            for (int e = 0; e < 4; e++)
                if (fv[e] > T)
                {
                    const int slot = atomicAdd(&bs.cnt, 1);
                    if (slot < kSurvCap)
                        bs.surv[slot] = make_int2(cc + e, __float_as_int(av[e]));
                }
        }
    }
    __syncthreads();
    const int cnt = bs.cnt;
    if (cnt < 2 || cnt > kSurvCap)
    {
        if (tid == 0)
            s.cache_hdr[i] = RowCacheHdr{0.0, 0, 0};
        __syncthreads();
        return false;
    }
    // ---- exact values, ranks -----------------------------------------------------------------------
    int2 mine = make_int2(0, 0);
    double d = 0.0;
    if (tid < cnt)
    {
        mine = bs.surv[tid];
        d = widen_exact(__int_as_float(mine.y)) - s.prices[mine.x];
        bs.sval[tid] = d;
    }
    __syncthreads();
    if (tid < cnt)
    {
        int rank = 0;
        for (int u = 0; u < cnt; u++)
        {
            const double y = bs.sval[u];
            rank += (y > d) || (y == d && bs.surv[u].x < mine.x);
        }
        if (rank < kCacheK)
            s.cache[(long long)i * kCacheK + rank] = mine;
        if (rank == 0)
        {
            bs.v = d;
            bs.j = mine.x;
            bs.a = __int_as_float(mine.y);
        }
        if (rank == 1)
            bs.w = d;
        if (rank == kCacheK)
            bs.kth = d;
    }
    if (tid >= cnt && tid < kCacheK)
        s.cache[(long long)i * kCacheK + tid] = make_int2(0, 0); // a == 0: not an entry
    __syncthreads();
    const double bound_T = (double)T + (double)margin;
    const bool ok = bs.w > bound_T;
    if (tid == 0)
    {
        RowCacheHdr h;
        h.bound = fmax(bound_T, bs.kth);
        h.count = ok ? min(cnt, kCacheK) : 0;
        h.pad = 0;
        s.cache_hdr[i] = h;
    }
    return ok;
}

// One warp: best / second best of row i from its cache and the current prices.  Returns true
// (all lanes) when the answer is exact; t then holds it in every lane and `prev` the current owner
// of the best column.  Unused cache slots hold a == 0 ("no edge", Auction.h:134), so the entry and
// the header load in parallel and the price gather waits for the entry alone.
struct CacheView
{
    const int2* cache;
    const RowCacheHdr* hdr;
    const double* prices;
    const int* owner;
};

__device__ __forceinline__ bool warp_cache_eval(const CacheView& c, int i, int lane, Top2<float>& t, int& prev)
{
    const int2 e = c.cache[(long long)i * kCacheK + lane];
    const RowCacheHdr h = c.hdr[i];
    const float a = __int_as_float(e.y);
    // prices and owners change every round: a line written by this SM is gone from L1 anyway
    // (stores invalidate), so read them where they live
    const double p = __ldcg(c.prices + e.x);
    const int own = __ldcg(c.owner + e.x);
    t = top2_empty<float>();
    if (a != 0.f && lane < h.count)
    {
        t.v = widen_exact(a) - p;
        t.j = e.x;
        t.a = a;
    }
    const int mine = t.j;
    top2_warp_reduce(t);
    const unsigned src = __ballot_sync(kFullMask, mine == t.j);
    prev = __shfl_sync(kFullMask, own, src ? __ffs(src) - 1 : 0);
    // w above the -1e6 start value is a real cached element (Auction.h:237); above `bound` it beats
    // every column outside the cache
    return h.count >= 2 && t.j >= 0 && t.w > h.bound && t.w > -kNoSecondBest;
}

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// All CTAs: a fresh cache for every row (entry of the narrow rounds of an eps phase).
__device__ __noinline__ void cache_build_sweep(const DenseState<float>& s, BuildShared& bs, float margin)
{
    const unsigned row_bytes = (unsigned)(s.cols4 * sizeof(float));
    const bool can_prefetch = (row_bytes & 15u) == 0 && ((s.ld * sizeof(float)) & 15u) == 0;
    for (int i = blockIdx.x; i < s.rows; i += gridDim.x)
    {
        if (can_prefetch && threadIdx.x == 0 && i + (int)gridDim.x < s.rows)
            prefetch_row_l2(s.A + (long long)(i + gridDim.x) * s.ld, row_bytes);
        cta_scan_build(s, bs, i, margin);
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// Controller CTA of the narrow rounds, cached form: finishes the eps phase that list[0..n) is the
// tail of (n <= kNarrowMax).  A round: one warp per bidder answers from the row cache; the rare
// misses are scanned by the whole CTA (which rebuilds their caches); the bids are resolved in
// shared memory with the rule of the wide rounds (highest bid per column, lowest row among equal
// bids; Auction.h:257-285 for the price / lock / evict update).  No other SM takes part: the
// scanner CTAs stay parked on their mailboxes until the phase is over.
// ---------------------------------------------------------------------------------------------
constexpr int kMissMark = -2;

// Every thread of the controller CTA: the bidders of the round whose cache could not answer
// (sh.col[b] == kMissMark) get a CTA scan of their row, which also rebuilds the cache.  Kept out of
// line: the scans need many more registers than the round loop around them.
__device__ __noinline__ void controller_misses(const DenseState<float>& s, NarrowShared<float>& sh, BuildShared& bs,
                                               const int* q, int n, double eps, double amax)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float margin = filter_margin(amax, sh.pmax);
    for (int b = 0; b < n; b++)
    {
        if (sh.col[b] != kMissMark) // the same for every thread
            continue;
        const int i = q[b];
        if (cta_scan_build(s, bs, i, margin))
        {
            if (tid == 0)
            {
                sh.col[b] = bs.j;
                sh.prev[b] = s.owner[bs.j];
                sh.bid[b] = (double)bs.a - bs.w + eps;
                s.profits[i] = bs.w - eps;
            }
        }
        else
        {
            // ties / too few edges: the unfiltered scan, with the -1e6 rescue (dense_kernels.cuh)
            if (tid == 0)
                sh.ctr[2]++;
            __syncthreads();
            Top2<float> m = cta_bid_scan(s, sh, i, margin, 0, s.cols4);
            if (warp == 0)
            {
                m = rescue_scan(s, i, m, lane);
                if (lane == 0)
                {
                    if (m.j < 0)
                    {
                        s.ctrl->status = 2; // B200LSAP_ERR_INFEASIBLE
                        sh.stop = 1;
                        sh.col[b] = -1;
                    }
                    else
                    {
                        sh.col[b] = m.j;
                        sh.prev[b] = s.owner[m.j];
                        sh.bid[b] = (double)m.a - m.w + eps;
                        s.profits[i] = m.w - eps;
                    }
                }
            }
        }
        __syncthreads();
    }
}

__device__ __noinline__ void narrow_controller_cached(const DenseState<float>& s, NarrowShared<float>& sh,
                                                         BuildShared& bs, const int* __restrict__ list, int n,
                                                         int mode_count, double eps, double amax, long long budget,
                                                         NarrowCounts& cnt)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int scanners = gridDim.x - 1;
    // entry: bidders in increasing row order (rank sort), so that the outcome does not depend on
    // the order in which the wide rounds' atomics compacted the list
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
        sh.nmiss2[0] = sh.nmiss2[1] = 0;
        sh.pmax = fmax(*reinterpret_cast<volatile double*>(&s.ctrl->pmax), amax);
    }
    __syncthreads();
    // counters live in shared memory (thread 0 only): at 64 registers per thread every 64-bit counter
    // kept in a register was spilled, and a spill slot re-read right after a store to its line
    // costs an L2 round trip per round
    if (tid < 4)
        sh.ctr[tid] = 0;
    if (tid == 0)
    {
        sh.inv.cache = s.cache;
        sh.inv.hdr = s.cache_hdr;
        sh.inv.prices = s.prices;
        sh.inv.pricesf = s.pricesf;
        sh.inv.profits = s.profits;
        sh.inv.owner = s.owner;
        sh.inv.row2col = s.row2col;
        sh.inv.eps = eps;
        sh.inv.budget = (int)min(budget, 0x7fffffffll);
        sh.round_no = 0;
    }
    __syncthreads();
    const volatile NarrowInvariants& inv = sh.inv;
#ifdef B200LSAP_TIMING
    if (tid < 16)
        sh.dbg[tid] = 0;
    __syncthreads();
#endif
    for (;;)
    {
        n = sh.qn;
        if (n == 0 || sh.stop)
            break;
        // rounds of this call so far: in shared memory like every other loop-carried value (a counter
        // in a register is spilled, and one spill store per round makes each reload of that stack
        // line an L2 round trip); its parity double-buffers the bidder list and the miss counter
        const int rounds = *reinterpret_cast<volatile int*>(&sh.round_no);
        if (rounds >= inv.budget)
        {
            if (tid == 0)
                s.ctrl->status = 3; // B200LSAP_ERR_MAX_ROUNDS
            break;
        }
        int* const q = (rounds & 1) ? sh.q2 : sh.q;  // bidders of this round
        int* const qn = (rounds & 1) ? sh.q : sh.q2; // the next round's
        TICK_INIT;
        // ---- bids from the caches: a warp per bidder ------------------------------------------------
        for (int b = warp; b < n; b += kWarpsPerBlock)
        {
            const int i = q[b];
            Top2<float> t;
            int prev;
            const CacheView cv{inv.cache, inv.hdr, inv.prices, inv.owner};
            const bool hit = warp_cache_eval(cv, i, lane, t, prev);
            TICK(10);
            if (lane == 0)
            {
                if (hit)
                {
                    sh.col[b] = t.j;
                    sh.prev[b] = prev;
                    const double e = inv.eps;
                    sh.bid[b] = (double)t.a - t.w + e; // Auction.h:250
                    inv.profits[i] = t.w - e;          // Auction.h:253
                }
                else
                {
                    sh.col[b] = kMissMark;
                    atomicAdd(&sh.nmiss2[rounds & 1], 1);
                }
            }
        }
        TICK(0);
        __syncthreads();
        TICK(1);
        // the counter of the next round is cleared now: its last readers passed the barrier that ended
        // the previous round, its next writers wait for the one that ends this round
        const int nmiss = sh.nmiss2[rounds & 1];
        if (tid == 0)
            sh.nmiss2[(rounds + 1) & 1] = 0;
        if (nmiss > 0)
        {
            controller_misses(s, sh, bs, q, n, inv.eps, amax);
            if (sh.stop)
                break;
        }
        TICK(2);
        if (tid == 0)
        {
            sh.ctr[0] += n;     // bids
            sh.ctr[1] += nmiss; // of which misses
        }
        // ---- resolve in shared memory: highest bid per column, lowest row among equal bids -------
        int my_row = -1, my_col = -1, my_prev = -1;
        double my_bid = 0.0;
        bool win = false;
        if (tid < n)
        {
            my_row = q[tid];
            my_col = sh.col[tid];
            my_bid = sh.bid[tid];
            win = true;
            for (int k = 0; k < n; k++)
                if (k != tid && sh.col[k] == my_col)
                {
                    const double ob = sh.bid[k];
                    const int orow = q[k];
                    if (ob > my_bid || (ob == my_bid && orow < my_row))
                        win = false;
                }
            TICK(6);
            if (win)
            {
                my_prev = sh.prev[tid];
                if (my_prev >= 0)
                {
                    // the evicted row bids next round: start pulling its cache towards this SM
#if B200LSAP_VAR != 2
                    const int2* const ce = inv.cache + (long long)my_prev * kCacheK;
                    prefetch_l1(ce);
                    prefetch_l1(ce + kCacheK / 2);
                    prefetch_l1(inv.hdr + my_prev);
#endif
                    inv.row2col[my_prev] = -1; // Auction.h:270
                }
                inv.prices[my_col] = my_bid; // Auction.h:259
                inv.pricesf[my_col] = (float)my_bid;
                inv.owner[my_col] = my_row;
                inv.row2col[my_row] = my_col;
                if (my_bid > sh.pmax)
                    atomicMax(reinterpret_cast<unsigned long long*>(&sh.pmax),
                              (unsigned long long)__double_as_longlong(my_bid));
            }
        }
        TICK(5);
        // ---- next list, in bidder order: a loser stays, a winner hands its place to the row it evicted
        const int push = tid < n ? (win ? my_prev : my_row) : -1;
        const unsigned mask = __ballot_sync(kFullMask, push >= 0);
        if (n <= 32)
        {
            if (warp == 0)
            {
                if (push >= 0)
                    qn[__popc(mask & ((1u << lane) - 1))] = push;
                if (lane == 0)
                    sh.qn = __popc(mask);
            }
        }
        else
        {
            if (lane == 0)
                sh.warp_count[warp] = __popc(mask);
            __syncthreads(); // warp_count is complete
            int base = 0, total = 0;
            for (int k = 0; k < kNarrowMax / 32; k++)
            {
                const int c = sh.warp_count[k];
                base += k < warp ? c : 0;
                total += c;
            }
            if (push >= 0)
                qn[base + __popc(mask & ((1u << lane) - 1))] = push;
            if (tid == 0)
                sh.qn = total;
        }
        if (tid == 0)
            sh.round_no = rounds + 1;
        TICK(3);
        __syncthreads();
        TICK(4);
    }
#ifdef B200LSAP_TIMING
    __syncthreads();
    if (tid < 8 || (tid >= 10 && tid < 16))
        s.ctrl->dbg[tid] += sh.dbg[tid];
    if (tid == 8)
        s.ctrl->dbg[8] += sh.round_no;
    if (tid == 9)
        s.ctrl->dbg[9] += sh.ctr[0];
#endif
    __syncthreads();
    cnt.rounds += sh.round_no;
    cnt.bids += sh.ctr[0];
    // release the scanners: a negative flag, different from the one of the previous narrow mode
    __syncthreads();
    if (tid < scanners)
    {
        __threadfence();
        *reinterpret_cast<volatile int*>(&s.narrow->slot[tid].flag) = -mode_count;
    }
    if (tid == 0)
    {
        if (sh.pmax > s.ctrl->pmax)
            s.ctrl->pmax = sh.pmax;
        s.ctrl->cache_hits += sh.ctr[0] - sh.ctr[1];
        s.ctrl->cache_misses += sh.ctr[1];
        s.ctrl->cache_fallbacks += sh.ctr[2];
    }
}

// ---------------------------------------------------------------------------------------------
// First eps of the scaling schedule, from the data.  Between two phases eps shrinks by a fixed
// factor; where to START is free (every phase but the last only prepares prices; the last runs at
// the reference's eps = 1e-5 / cols, Auction.h:55, which alone decides the result).  A phase whose
// eps exceeds the spread of a row's best few dozen values only rotates prices: every bid pushes
// its column below all the others.  k_probe measures that spread on kProbeRows sampled rows at the
// initial prices: best value minus the cache bound (about the 36th best); k_schedule starts at a
// quarter of the mean.  Rows the CTA scan cannot answer (ties) do not count; with fewer than half
// of the samples usable the classic start of 1/4 of the coefficient range is kept.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1) k_probe(DenseState<float> s)
{
    __shared__ BuildShared bs;
    const double amax = s.ctrl->amax;
    const float margin = filter_margin(amax, fmax(s.ctrl->pmax, amax));
    const int i = (int)((long long)blockIdx.x * s.rows / gridDim.x);
    const bool ok = cta_scan_build(s, bs, i, margin);
    if (threadIdx.x == 0)
        s.ctrl->probe[blockIdx.x] = ok ? bs.v - s.cache_hdr[i].bound : -1.0;
}

__global__ void k_schedule(DenseCtrl* ctrl, int n_probe)
{
    if (threadIdx.x != 0 || blockIdx.x != 0 || !ctrl->eps_auto)
        return;
    double sum = 0.0;
    int used = 0;
    for (int k = 0; k < n_probe; k++)
        if (ctrl->probe[k] >= 0.0)
        {
            sum += ctrl->probe[k];
            used++;
        }
    double e = 0.25;
    if (used * 2 >= n_probe && used > 0)
        e = fmin(0.25, sum / used * 0.25);
    int n = 0;
    while (e > ctrl->eps_final * 1.0000001 && n < kMaxPhases - 1)
    {
        ctrl->eps[n++] = e;
        e /= ctrl->eps_factor;
    }
    ctrl->eps[n++] = ctrl->eps_final;
    ctrl->n_phases = n;
}

} // namespace b200lsap
