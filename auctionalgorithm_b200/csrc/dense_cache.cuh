// dense_cache.cuh — row caches: narrow rounds without row scans (fp32 and fp64 coefficients; a cache
// entry is (column, a): 8 bytes for float, 16 for double).
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

template <typename T> struct BuildShared
{
    float cand[64];             // group g: cand[g] = largest f of its elements, cand[32 + g] = second largest
    CacheEnt<T> surv[kSurvCap]; // survivors: (column, a)
    double sval[kSurvCap];      // their exact values
    int cnt;
    int ok;
    float thr; // the selection threshold
    int j;
    T a;
    double v, w, kth;
};

// pieces of a row a thread keeps in registers between the two passes: 16384 columns for either CTA size
template <int kT> struct BuildRegPieces
{
    static constexpr int value = 4096 / kT;
};

// f = fl32(a) - fl32(p) of one 4-element piece; a == 0 is "no edge".  For fp64 coefficients the
// conversion adds at most 2^-24 |a| to the filter's error, still well inside filter_margin.
template <typename V4> __device__ __forceinline__ float4 piece_f(const V4 a, const float4 q)
{
    return make_float4(a.x != 0 ? (float)a.x - q.x : -INFINITY, a.y != 0 ? (float)a.y - q.y : -INFINITY,
                       a.z != 0 ? (float)a.z - q.z : -INFINITY, a.w != 0 ? (float)a.w - q.w : -INFINITY);
}

template <typename T> __device__ __forceinline__ typename Vec4<T>::type vec4_zero()
{
    typename Vec4<T>::type v;
    v.x = v.y = v.z = v.w = T(0);
    return v;
}

// Every thread of the CTA.  Scans row i, rebuilds its cache, returns true when the result
// (bs.v, bs.w, bs.j, bs.a) is exact; all threads see the same return value and may read bs until
// the next call.  On false the row's cache is left empty.
// kT = threads that share the row: the whole CTA (k_probe 1024, k_narrow 512; tid = threadIdx.x, barrier
// 0), or a quarter of the persistent kernel's CTA (cache_build_sweep: 256 threads with a named barrier
// of their own, four rows in flight per SM).  The 64 threshold candidates are the top-2 of 32 groups
// of kT / 32 consecutive threads.  kBuildRegPieces pieces of the row stay in registers between the
// two passes; the rest (all of it with 0) is read again, from L2.
template <int kT> __device__ __forceinline__ void group_bar(int bar_id)
{
    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(kT) : "memory");
}

template <int kT, int kBuildRegPieces, typename T>
__device__ __forceinline__ bool cta_scan_build(const DenseState<T>& s, BuildShared<T>& bs, int i, float margin,
                                               int tid = threadIdx.x, int bar_id = 0)
{
    using V4 = typename Vec4<T>::type;
    constexpr int kGroup = kT / 32; // lanes per candidate group
    const int lane = tid & 31;
    const T* __restrict__ row = s.A + (long long)i * s.ld;
    const int c1 = s.cols4;
    constexpr int stride = 4 * kT;
    V4 areg[kBuildRegPieces > 0 ? kBuildRegPieces : 1];
    // ---- pass 1: f = a - (float)p, top-2 of the thread's elements -------------------------------
    float hi = -INFINITY, lo = -INFINITY;
#pragma unroll
    for (int u = 0; u < kBuildRegPieces; u++)
    {
        const int cc = tid * 4 + u * stride;
        areg[u] = vec4_zero<T>();
        if (cc < c1)
            areg[u] = ld_stream4_evict_first(row + cc);
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
    for (int c0 = tid * 4 + kBuildRegPieces * stride; c0 < c1; c0 += 4 * stride) // what registers do not hold
    {
        V4 a4[4];
#pragma unroll
        for (int u = 0; u < 4; u++)
        {
            a4[u] = vec4_zero<T>();
            if (c0 + u * stride < c1)
                a4[u] = ld_stream4(row + c0 + u * stride);
        }
#pragma unroll
        for (int u = 0; u < 4; u++)
        {
            const int cc = c0 + u * stride;
            if (cc < c1)
            {
                const float4 f = piece_f(a4[u], *reinterpret_cast<const float4*>(s.pricesf + cc));
                lo = fmaxf(lo, fminf(hi, f.x)), hi = fmaxf(hi, f.x);
                lo = fmaxf(lo, fminf(hi, f.y)), hi = fmaxf(hi, f.y);
                lo = fmaxf(lo, fminf(hi, f.z)), hi = fmaxf(hi, f.z);
                lo = fmaxf(lo, fminf(hi, f.w)), hi = fmaxf(hi, f.w);
            }
        }
    }
#pragma unroll
    for (int sft = kGroup / 2; sft >= 1; sft >>= 1)
    {
        const float ohi = __shfl_xor_sync(kFullMask, hi, sft);
        const float olo = __shfl_xor_sync(kFullMask, lo, sft);
        lo = fmaxf(fminf(hi, ohi), fmaxf(lo, olo));
        hi = fmaxf(hi, ohi);
    }
    if ((lane & (kGroup - 1)) == 0)
    {
        bs.cand[tid / kGroup] = hi;
        bs.cand[32 + tid / kGroup] = lo;
    }
    if (tid == 0)
    {
        bs.cnt = 0;
        bs.ok = 0;
        bs.thr = INFINITY;
        bs.kth = -CUDART_INF;
    }
    group_bar<kT>(bar_id);
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
            bs.thr = x - 2.f * margin;
    }
    group_bar<kT>(bar_id);
    const float thr = bs.thr;
    if (!(thr < INFINITY)) // fewer than two edges in the row
    {
        if (tid == 0)
            s.cache_hdr[i] = RowCacheHdr{0.0, 0, 0};
        group_bar<kT>(bar_id);
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
            if (fmaxf(fmaxf(f.x, f.y), fmaxf(f.z, f.w)) > thr)
            {
                const float fv[4] = {f.x, f.y, f.z, f.w};
                const T av[4] = {areg[u].x, areg[u].y, areg[u].z, areg[u].w};
#pragma unroll
                for (int e = 0; e < 4; e++)
                    if (fv[e] > thr)
                    {
                        const int slot = atomicAdd(&bs.cnt, 1);
                        if (slot < kSurvCap)
                            bs.surv[slot] = make_ent<T>(cc + e, av[e]);
                    }
            }
        }
    }
    for (int c0 = tid * 4 + kBuildRegPieces * stride; c0 < c1; c0 += 4 * stride) // re-read from L2, last use
    {
        V4 a4[4];
#pragma unroll
        for (int u = 0; u < 4; u++)
        {
            a4[u] = vec4_zero<T>();
            if (c0 + u * stride < c1)
                a4[u] = ld_stream4_evict_first(row + c0 + u * stride);
        }
#pragma unroll
        for (int u = 0; u < 4; u++)
        {
            const int cc = c0 + u * stride;
            if (cc < c1)
            {
                const float4 f = piece_f(a4[u], *reinterpret_cast<const float4*>(s.pricesf + cc));
                if (fmaxf(fmaxf(f.x, f.y), fmaxf(f.z, f.w)) > thr)
                {
                    const float fv[4] = {f.x, f.y, f.z, f.w};
                    const T av[4] = {a4[u].x, a4[u].y, a4[u].z, a4[u].w};
#pragma unroll
                    for (int e = 0; e < 4; e++)
                        if (fv[e] > thr)
                        {
                            const int slot = atomicAdd(&bs.cnt, 1);
                            if (slot < kSurvCap)
                                bs.surv[slot] = make_ent<T>(cc + e, av[e]);
                        }
                }
            }
        }
    }
    group_bar<kT>(bar_id);
    const int cnt = bs.cnt;
    if (cnt < 2 || cnt > kSurvCap)
    {
        if (tid == 0)
            s.cache_hdr[i] = RowCacheHdr{0.0, 0, 0};
        group_bar<kT>(bar_id);
        return false;
    }
    // ---- exact values, ranks -----------------------------------------------------------------------
    CacheEnt<T> mine = make_ent<T>(0, T(0));
    double d = 0.0;
    if (tid < cnt)
    {
        mine = bs.surv[tid];
        d = widen_exact(mine.a) - s.prices[mine.col];
        bs.sval[tid] = d;
    }
    group_bar<kT>(bar_id);
    if (tid < cnt)
    {
        int rank = 0;
        for (int u = 0; u < cnt; u++)
        {
            const double y = bs.sval[u];
            rank += (y > d) || (y == d && bs.surv[u].col < mine.col);
        }
        if (rank < kCacheK)
            s.cache[(long long)i * kCacheK + rank] = mine;
        if (rank == 0)
        {
            bs.v = d;
            bs.j = mine.col;
            bs.a = mine.a;
        }
        if (rank == 1)
            bs.w = d;
        if (rank == kCacheK)
            bs.kth = d;
    }
    if (tid >= cnt && tid < kCacheK)
        s.cache[(long long)i * kCacheK + tid] = make_ent<T>(0, T(0)); // a == 0: not an entry
    group_bar<kT>(bar_id);
    const double bound_T = (double)thr + (double)margin;
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

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// All CTAs: a fresh cache for every row (entry of the narrow rounds of an eps phase).  A scan is a
// chain of latencies (HBM, six barriers, two small rank loops: ~8 us when a whole CTA takes one row at
// a time), so every CTA works on four rows at once: 256 threads and a named barrier per row, the
// row read twice (HBM, then L2) instead of being held in registers.
template <typename T> struct SweepGroups
{
    // 8 rows in flight per CTA for fp32; fp64 survivors take twice the shared memory: 4 rows
    static constexpr int threads = sizeof(T) == 4 ? 128 : 256;
    static constexpr int count = kThreads / threads;
};

template <typename T>
__device__ __noinline__ void cache_build_sweep(const DenseState<T>& s, BuildShared<T>* bs, float margin)
{
    constexpr int kGT = SweepGroups<T>::threads, kG = SweepGroups<T>::count;
    const int g = threadIdx.x / kGT, gtid = threadIdx.x % kGT;
    for (int i = blockIdx.x * kG + g; i < s.rows; i += gridDim.x * kG)
    {
        cta_scan_build<kGT, 0, T>(s, bs[g], i, margin, gtid, 1 + g);
        group_bar<kGT>(1 + g); // bs[g] is reused by the next row
    }
}

// ---------------------------------------------------------------------------------------------
// First eps of the scaling schedule, from the data.  Between two phases eps shrinks by a fixed
// factor; where to START is free (every phase but the last only prepares prices; the last runs at
// the reference's eps = 1e-5 / cols, Auction.h:55, which alone decides the result).  A phase whose
// eps exceeds the spread of a row's best few dozen values only rotates prices: every bid pushes
// its column below all the others.  k_probe measures that spread on kProbeRows sampled rows at the
// initial prices: best value minus the cache bound (about the 36th best); k_schedule starts at an
// eighth of the mean (measured on B200 over sizes 4096-16384 and three seeds: start 1/8 or 1/4 of the
// spread and factor 7 gave the shortest solves; the landscape is noisy, +-20 % between seeds).  Rows the CTA scan cannot answer (ties) do not count; with fewer than half
// of the samples usable the classic start of 1/4 of the coefficient range is kept.
// ---------------------------------------------------------------------------------------------
template <typename T> __global__ void __launch_bounds__(kThreads, 1) k_probe(DenseState<T> s)
{
    __shared__ BuildShared<T> bs;
    const double amax = s.ctrl->amax;
    const float margin = filter_margin(amax, fmax(s.ctrl->pmax, amax));
    const int i = (int)((long long)blockIdx.x * s.rows / gridDim.x);
    // fp64 rows are not held in registers between the passes (they would need twice as many)
    const bool ok = cta_scan_build<kThreads, (sizeof(T) == 4 ? BuildRegPieces<kThreads>::value : 0), T>(s, bs, i, margin);
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
        e = fmin(0.25, sum / used * ctrl->eps_probe_scale);
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
