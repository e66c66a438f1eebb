// narrow_kernel.cuh — k_narrow<kSmem, T>: the narrow rounds of an eps phase on ONE SM.
//
// Included by dense_kernels.cuh after dense_cache.cuh.
//
// Once a phase is down to <= kNarrowMax bidders the cooperative kernel stops (after giving every row
// a fresh cache, dense_cache.cuh) and this single-CTA kernel finishes the phase; the host has both
// launches of every phase enqueued beforehand and is not touched per round.  Why a kernel of its own:
//   * a round is a chain of dependent latencies, not bandwidth.  Measured on B200 (tools/micro):
//     L1 hit 39 cycles, L2 hit ~290, a load of a line this SM has just stored to ~660 (stores
//     invalidate the L1 line and the load queues behind the store), shared memory ~35, bar.sync
//     ~50-80.  Prices and owners change every round, so inside the cooperative kernel each round
//     paid two or three L2 round trips for them;
//   * here prices[] and owner[] live in SHARED MEMORY for the whole phase (12 bytes per column:
//     16384 columns = 192 KB of the SM's 227 KB; kSmem, chosen by the host when they fit).  Global
//     memory is not touched per round: prices are flushed before any row scan (narrow_flush_prices)
//     and everything is written back once at the end (narrow_writeback rebuilds row2col from
//     owner).  The cooperative kernel cannot afford this: its sweeps need that space as L1;
//   * 512 threads at up to 128 registers instead of 1024 at 64: the round loop holds its state in
//     registers (at 64 every loop-carried value was spilled, and a spill store per round turns the
//     reloads from that stack line into L2 round trips);
//   * the top-2 of the 32 cached values is five redux.sync on the ordered bits of the doubles
//     instead of five levels of 64-bit shuffles and dependent fp64 compares (~200 vs ~550 cycles);
//   * a round with ONE bidder (most rounds of a square problem) never leaves warp 0: evaluate, bid,
//     evict, continue with the evicted row, no barrier and no shared-memory exchange.
// The rounds themselves are unchanged (Jacobi rounds of the n unassigned rows; highest bid per
// column, lowest row among equal bids; price / lock / evict as Auction.h:250-285), so prices come out
// bit-identical to the row-scanning paths (tests: test_golden_layouts_dtypes_and_drivers_agree).
#pragma once

namespace b200lsap
{

constexpr int kNarrowThreads = 512;
constexpr int kNarrowWarps = kNarrowThreads / 32;
constexpr int kMissMark = -2;
#ifndef B200LSAP_SOLO_MAX
#define B200LSAP_SOLO_MAX 2
#endif
constexpr int kSoloMax = B200LSAP_SOLO_MAX; // rounds with 2..kSoloMax bidders are played by warp 0 alone
constexpr int kNarrowHashBits = 9, kNarrowHash = 1 << kNarrowHashBits; // >= 2 kNarrowMax slots

template <typename T> struct NarrowKShared
{
    BuildShared<T> bs;
    Top2<T> part[kNarrowWarps]; // unfiltered fallback scan
    int q[2][kNarrowMax];           // bidders of the round / of the next round
    int col[kNarrowMax];            // column each bidder bid for (kMissMark: cache could not answer)
    int prev[kNarrowMax];           // current owner of that column
    double bid[kNarrowMax];
    int warp_count[kNarrowWarps];
    int hcol[kNarrowHash];  // resolve of a round with more than 32 bidders: column of the slot, -1 empty
    int hbest[kNarrowHash]; //   bidder index of the best bid for it so far, -1 none
    int qn, stop, nmiss, served, served_bids;
    double pmax;
    Top2<T> res; // result of the fallback scan
};

// Order-preserving map of a double to u64 (larger double <=> larger key); -0.0 is folded into +0.0
// first so that equal values have equal keys, as the reference's '>' sees them.
__device__ __forceinline__ unsigned long long ordered_key(double v)
{
    const long long b = __double_as_longlong(v + 0.0);
    return (unsigned long long)(b ^ ((b >> 63) | (long long)0x8000000000000000ull));
}
__device__ __forceinline__ double key_to_double(unsigned long long k)
{
    const long long b = (long long)k;
    return __longlong_as_double(b < 0 ? (b ^ (long long)0x8000000000000000ull) : ~b);
}

// One warp: best / second best of row i from its cache (dense_cache.cuh) and the current prices.
// Returns true (all lanes) when the answer is exact; t then holds it in every lane and `prev` the
// current owner of the best column.  Unused cache slots hold a == 0 ("no edge", Auction.h:134).
// Tie rules: the best is the largest value, lowest column among equals (strict '>' in column
// order, Auction.h:136); the second best may equal the best (Auction.h:147-148).
#ifdef B200LSAP_TIMING
__device__ long long g_ev[8]; // developer counters of the one-bidder path
#endif

template <bool kSmem, bool kChase, typename T>
__device__ __forceinline__ bool narrow_eval_regs(const DenseState<T>& s, const double* prices_s, const int* owner_s,
                                                 const CacheEnt<T> e, const RowCacheHdr h, int lane, Top2<T>& t, int& prev,
                                                 long long* evp);

template <bool kSmem, bool kChase = false, typename T>
__device__ __forceinline__ bool narrow_eval(const DenseState<T>& s, const double* prices_s, const int* owner_s,
                                            int i, int lane, Top2<T>& t, int& prev, long long* evp = nullptr)
{
    return narrow_eval_regs<kSmem, kChase, T>(s, prices_s, owner_s, s.cache[(long long)i * kCacheK + lane], s.cache_hdr[i],
                                           lane, t, prev, evp);
}

// The same with the row's cache entry and header already loaded (one-bidder chain: the next
// row's loads are issued before the current bid is committed).
template <bool kSmem, bool kChase, typename T>
__device__ __forceinline__ bool narrow_eval_regs(const DenseState<T>& s, const double* prices_s, const int* owner_s,
                                                 const CacheEnt<T> e, const RowCacheHdr h, int lane, Top2<T>& t, int& prev,
                                                 long long* evp)
{
#ifdef B200LSAP_TIMING
    long long ev0 = clock64();
#define EVTICK(k, ...)                                                                                             \
    do                                                                                                             \
    {                                                                                                              \
        if (kChase)                                                                                                \
        {                                                                                                          \
            asm volatile("" ::__VA_ARGS__ : "memory");                                                             \
            const long long now_ = clock64();                                                                      \
            evp[k] += now_ - ev0;                                                                                  \
            ev0 = now_;                                                                                            \
        }                                                                                                          \
    } while (0)
#else
#define EVTICK(k, ...)
#endif
    EVTICK(0, "r"(e.col), "r"(h.count));
    const T a = e.a;
    double p;
    int own;
    if (kSmem)
    {
        p = prices_s[e.col];
        own = owner_s[e.col];
    }
    else
    {
        p = __ldcg(s.prices + e.col);
        own = __ldcg(s.owner + e.col);
    }
    EVTICK(1, "d"(p), "r"(own));
    const bool valid = a != T(0) && lane < h.count;
    const unsigned long long key = valid ? ordered_key(widen_exact(a) - p) : 0ull;
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    // best: max of the high words, then of the low words among those, then the lowest column
    const unsigned mh = __reduce_max_sync(kFullMask, hi);
    const bool c1 = valid && hi == mh;
    const unsigned ml = __reduce_max_sync(kFullMask, c1 ? lo : 0u);
    const bool top = c1 && lo == ml;
    const int mj = (int)__reduce_min_sync(kFullMask, top ? (unsigned)e.col : 0xffffffffu);
    const bool winner = top && e.col == mj;
    // second best: the same without the winner
    const unsigned hi2 = winner ? 0u : hi;
    const unsigned mh2 = __reduce_max_sync(kFullMask, hi2);
    const unsigned ml2 = __reduce_max_sync(kFullMask, (!winner && valid && hi == mh2) ? lo : 0u);
    EVTICK(2, "r"(mh2), "r"(ml2), "r"(mj));
    const unsigned wmask = __ballot_sync(kFullMask, winner);
    const int src = wmask ? __ffs(wmask) - 1 : 0;
    t.a = __shfl_sync(kFullMask, a, src);
    prev = __shfl_sync(kFullMask, own, src);
    t.j = wmask ? mj : -1;
    t.v = key_to_double(((unsigned long long)mh << 32) | ml);
    const unsigned long long k2 = ((unsigned long long)mh2 << 32) | ml2;
    t.w = k2 ? key_to_double(k2) : -kNoSecondBest;
    EVTICK(3, "r"(prev), "d"(t.w));
    // w above the -1e6 start value of the scan (Auction.h:237) and above `bound` beats every
    // column outside the cache
    return h.count >= 2 && wmask != 0 && k2 != 0 && t.w > h.bound && t.w > -kNoSecondBest;
}

// Two bidders of one warp at once (rounds with more bidders than warps): the same steps as
// narrow_eval_regs with the two dependency chains interleaved stage by stage, so that the latency of
// one bidder's load, gather or reduction hides behind the other's.  Results as narrow_eval.
template <bool kSmem, typename T>
__device__ __forceinline__ void narrow_eval2(const DenseState<T>& s, const double* prices_s, const int* owner_s, const int (&i)[2],
                                             int lane, Top2<T> (&t)[2], int (&prev)[2], bool (&hit)[2])
{
    CacheEnt<T> e[2];
    RowCacheHdr h[2];
#pragma unroll
    for (int u = 0; u < 2; u++)
    {
        e[u] = s.cache[(long long)i[u] * kCacheK + lane];
        h[u] = s.cache_hdr[i[u]];
    }
    double p[2];
    int own[2];
#pragma unroll
    for (int u = 0; u < 2; u++)
    {
        if (kSmem)
        {
            p[u] = prices_s[e[u].col];
            own[u] = owner_s[e[u].col];
        }
        else
        {
            p[u] = __ldcg(s.prices + e[u].col);
            own[u] = __ldcg(s.owner + e[u].col);
        }
    }
    bool valid[2], c1[2], top[2], winner[2];
    unsigned hi[2], lo[2], mh[2], ml[2], mh2[2], ml2[2], wmask[2];
    int mj[2];
#pragma unroll
    for (int u = 0; u < 2; u++)
    {
        valid[u] = e[u].a != T(0) && lane < h[u].count;
        const unsigned long long key = valid[u] ? ordered_key(widen_exact(e[u].a) - p[u]) : 0ull;
        hi[u] = (unsigned)(key >> 32);
        lo[u] = (unsigned)key;
    }
#pragma unroll
    for (int u = 0; u < 2; u++)
        mh[u] = __reduce_max_sync(kFullMask, hi[u]);
#pragma unroll
    for (int u = 0; u < 2; u++)
    {
        c1[u] = valid[u] && hi[u] == mh[u];
        ml[u] = __reduce_max_sync(kFullMask, c1[u] ? lo[u] : 0u);
    }
#pragma unroll
    for (int u = 0; u < 2; u++)
    {
        top[u] = c1[u] && lo[u] == ml[u];
        mj[u] = (int)__reduce_min_sync(kFullMask, top[u] ? (unsigned)e[u].col : 0xffffffffu);
    }
#pragma unroll
    for (int u = 0; u < 2; u++)
    {
        winner[u] = top[u] && e[u].col == mj[u];
        mh2[u] = __reduce_max_sync(kFullMask, winner[u] ? 0u : hi[u]);
    }
#pragma unroll
    for (int u = 0; u < 2; u++)
        ml2[u] = __reduce_max_sync(kFullMask, (!winner[u] && valid[u] && hi[u] == mh2[u]) ? lo[u] : 0u);
#pragma unroll
    for (int u = 0; u < 2; u++)
        wmask[u] = __ballot_sync(kFullMask, winner[u]);
#pragma unroll
    for (int u = 0; u < 2; u++)
    {
        const int src = wmask[u] ? __ffs(wmask[u]) - 1 : 0;
        t[u].a = __shfl_sync(kFullMask, e[u].a, src);
        prev[u] = __shfl_sync(kFullMask, own[u], src);
        t[u].j = wmask[u] ? mj[u] : -1;
        t[u].v = key_to_double(((unsigned long long)mh[u] << 32) | ml[u]);
        const unsigned long long k2 = ((unsigned long long)mh2[u] << 32) | ml2[u];
        t[u].w = k2 ? key_to_double(k2) : -kNoSecondBest;
        hit[u] = h[u].count >= 2 && wmask[u] != 0 && k2 != 0 && t[u].w > h[u].bound && t[u].w > -kNoSecondBest;
    }
}

// Every thread: unfiltered scan of row i (ties, too few edges), with the -1e6 rescue of rescue_scan.
// Result in sh.res, valid after the trailing barrier.
template <typename T> __device__ __noinline__ void narrow_exact_scan(const DenseState<T>& s, NarrowKShared<T>& sh, int i)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const T* __restrict__ row = s.A + (long long)i * s.ld;
    Top2<T> t = top2_empty<T>();
    for (int c = tid * 4; c < s.cols4; c += 4 * kNarrowThreads)
    {
        const auto a = ld_stream4(row + c);
        top2_push(t, a.x, s.prices[c], c);
        top2_push(t, a.y, s.prices[c + 1], c + 1);
        top2_push(t, a.z, s.prices[c + 2], c + 2);
        top2_push(t, a.w, s.prices[c + 3], c + 3);
    }
    top2_warp_reduce(t);
    if (lane == 0)
        sh.part[warp] = t;
    __syncthreads();
    if (warp == 0)
    {
        Top2<T> m = lane < kNarrowWarps ? sh.part[lane] : top2_empty<T>();
        top2_warp_reduce(m);
        m = rescue_scan(s, i, m, lane);
        if (lane == 0)
            sh.res = m;
    }
    __syncthreads();
}

// Every thread: bidders whose cache could not answer get a CTA scan (which rebuilds the cache).
template <bool kSmem, typename T>
__device__ __noinline__ void narrow_misses(const DenseState<T>& s, NarrowKShared<T>& sh, const int* owner_s, const int* q,
                                           int n, double eps, double amax)
{
    const int tid = threadIdx.x;
    const float margin = filter_margin(amax, sh.pmax);
    for (int b = 0; b < n; b++)
    {
        if (sh.col[b] != kMissMark) // the same for every thread
            continue;
        const int i = q[b];
        int j = -1;
        T a = T(0);
        double w = 0.0;
        // fp64 rows are not held in registers between the passes (they would need twice as many)
        if (cta_scan_build<kNarrowThreads, (sizeof(T) == 4 ? BuildRegPieces<kNarrowThreads>::value : 0), T>(s, sh.bs, i, margin))
        {
            j = sh.bs.j;
            a = sh.bs.a;
            w = sh.bs.w;
        }
        else
        {
            if (tid == 0)
                s.ctrl->cache_fallbacks++;
            __syncthreads();
            narrow_exact_scan(s, sh, i);
            j = sh.res.j;
            a = sh.res.a;
            w = sh.res.w;
        }
        if (tid == 0)
        {
            if (j < 0)
            {
                s.ctrl->status = 2; // B200LSAP_ERR_INFEASIBLE: the reference would never return
                sh.stop = 1;
                sh.col[b] = -1;
            }
            else
            {
                sh.col[b] = j;
                sh.prev[b] = kSmem ? owner_s[j] : s.owner[j];
                sh.bid[b] = (double)a - w + eps; // Auction.h:250
                s.profits[i] = w - eps;          // Auction.h:253
            }
        }
        __syncthreads();
    }
}

// price / lock / evict of one winning bid (Auction.h:259-270).  With prices and owners in shared
// memory nothing goes to global memory here: the same few lines would be stored to round after round
// (a price war hits the same columns), and those stores cost the general round ~340 cycles (measured).
// narrow_writeback publishes the state at the end of the kernel, and before any row scan.
template <bool kSmem, typename T>
__device__ __forceinline__ void narrow_commit(const DenseState<T>& s, double* prices_s, int* owner_s, int row, int col,
                                              double bid, int prev)
{
    if (kSmem)
    {
        prices_s[col] = bid;
        owner_s[col] = row;
    }
    else
    {
        s.prices[col] = bid;
        s.pricesf[col] = (float)bid;
        s.owner[col] = row;
        s.row2col[row] = col;
        if (prev >= 0)
            s.row2col[prev] = -1;
    }
}

// Every thread: global prices / pricesf / owner from the shared-memory copies (barrier included).
template <typename T>
__device__ __forceinline__ void narrow_flush_prices(const DenseState<T>& s, const double* prices_s, const int* owner_s)
{
    for (int j = threadIdx.x; j < s.cols; j += kNarrowThreads)
    {
        const double p = prices_s[j];
        s.prices[j] = p;
        s.pricesf[j] = (float)p;
        s.owner[j] = owner_s[j];
    }
    __syncthreads();
}

// Every thread, end of the kernel: prices, owners, and row2col rebuilt as the inverse of owner.
template <typename T>
__device__ __forceinline__ void narrow_writeback(const DenseState<T>& s, const double* prices_s, const int* owner_s)
{
    for (int i = threadIdx.x; i < s.rows; i += kNarrowThreads)
        s.row2col[i] = -1;
    narrow_flush_prices(s, prices_s, owner_s);
    for (int j = threadIdx.x; j < s.cols; j += kNarrowThreads)
    {
        const int o = owner_s[j];
        if (o >= 0)
            s.row2col[o] = j;
    }
}

// Safeguard of the data-sized first eps (k_probe / k_schedule).  The probe sizes eps[0] by the spread of a
// row's best values; on inputs whose equilibrium prices are far larger than that spread (every row
// wants the same few columns: a = u_i * v_j; or rows with very few edges) the first phase would crawl
// there in steps of eps.  The first phase of the schedule therefore has a work budget (rounds + 4 per
// cache miss, a few times the number of rows); when it is spent the schedule restarts from
// eps * factor^2 (at most eps_cap, the classic start), prices kept, and that phase gets a budget again.
// Only the last phase, at the reference's eps, decides the result, so any such restart is legal.
__device__ __forceinline__ void escalate_schedule(DenseCtrl* ctrl, int cur)
{
    const int k = ctrl->phase;
    const double f = ctrl->eps_factor;
    double e = fmin(ctrl->eps_cap, ctrl->eps[k] * f * f);
    int n = k;
    while (e > ctrl->eps_final * 1.0000001 && n < kMaxPhases - 1)
    {
        ctrl->eps[n++] = e;
        e /= f;
    }
    ctrl->eps[n++] = ctrl->eps_final;
    ctrl->n_phases = n;
    ctrl->esc_phase = k;
    ctrl->phase = k - 1; // the next cooperative launch advances to phase k and resets the assignment
    ctrl->count[cur] = 0;
    ctrl->phase_work = 0;
    ctrl->escalations++;
}

template <bool kSmem, typename T>
__global__ void __launch_bounds__(kNarrowThreads, 1) k_narrow(const __grid_constant__ DenseState<T> s)
{
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    __shared__ NarrowKShared<T> sh;
    double* const prices_s = reinterpret_cast<double*>(dyn_smem);
    int* const owner_s = reinterpret_cast<int*>(prices_s + ((s.cols + 1) & ~1));
    DenseCtrl* const ctrl = s.ctrl;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (ctrl->done || ctrl->status != 0)
        return;
    const int cur = ctrl->cur;
    int n = ctrl->count[cur];
    if (n <= 0 || n > kNarrowMax)
        return;
    const double eps = ctrl->eps[ctrl->phase];
    const double amax = ctrl->amax;
    const long long budget_ll = ctrl->max_rounds - ctrl->rounds;
    const int budget = (int)max(0ll, min(budget_ll, 0x7fffffffll));
#ifdef B200LSAP_TIMING
    const long long t_begin = clock64();
#endif
    if (kSmem)
        for (int j = tid; j < s.cols; j += kNarrowThreads)
        {
            prices_s[j] = s.prices[j];
            owner_s[j] = s.owner[j];
        }
#if B200LSAP_VAR == 3
    {
        // experiment: pull every row cache into L2
        int acc = 0;
        const int4* c4 = reinterpret_cast<const int4*>(s.cache);
        const long long n4 = (long long)s.rows * kCacheK / 2;
        for (long long k = tid; k < n4; k += kNarrowThreads)
            acc += __ldcg(c4 + k).x;
        const int4* h4 = reinterpret_cast<const int4*>(s.cache_hdr);
        for (long long k = tid; k < s.rows; k += kNarrowThreads)
            acc += __ldcg(h4 + k).z;
        if (acc == 0x7fffffff)
            sh.stop = 1;
    }
#endif
    // bidders in increasing row order (rank sort), so that the outcome does not depend on the order
    // in which the wide rounds' atomics compacted the list
    const int* __restrict__ list = s.list[cur];
    if (tid < n)
    {
        const int mine = list[tid];
        int rank = 0;
        for (int k = 0; k < n; k++)
            rank += list[k] < mine;
        sh.q[0][rank] = mine;
    }
    for (int k = tid; k < kNarrowHash; k += kNarrowThreads)
    {
        sh.hcol[k] = -1;
        sh.hbest[k] = -1;
    }
    if (tid == 0)
    {
        sh.qn = n;
        sh.stop = 0;
        sh.nmiss = 0;
        sh.served = 0;
        sh.pmax = fmax(ctrl->pmax, amax);
    }
    __syncthreads();
    int rounds = 0, buf = 0;
    const bool esc_ok = ctrl->phase_budget > 0 && ctrl->phase == ctrl->esc_phase && eps < ctrl->eps_cap;
    const long long work_budget = ctrl->phase_budget;
    long long work = ctrl->phase_work; // the wide rounds of this phase
    bool escalated = false;
    long long bids = 0, misses = 0, fast = 0, solo_rounds = 0;
    // (an entry read before its row's cache was rebuilt stays valid: bounds hold for ever, prices only rise)
    int pre_row = -1;
    CacheEnt<T> pre_e = make_ent<T>(0, T(0));
    RowCacheHdr pre_h{0.0, 0, 0};
#ifdef B200LSAP_TIMING
    long long acc2 = 0, acc3 = 0, acc4 = 0, acc5 = 0, acc6 = 0, acc7 = 0;
    long long bk_c0 = 0, bk_c1 = 0, bk_c2 = 0, bk_c3 = 0, bk_r0 = 0, bk_r1 = 0, bk_r2 = 0, bk_r3 = 0;
#endif
    for (;;)
    {
        n = sh.qn;
        if (n == 0 || sh.stop)
            break;
        if (rounds >= budget)
        {
            if (tid == 0)
                ctrl->status = 3; // B200LSAP_ERR_MAX_ROUNDS
            break;
        }
        if (esc_ok && work > work_budget)
        {
            escalated = true; // the same for every thread
            break;
        }
        int* const q = sh.q[buf];
        int* const qn = sh.q[buf ^ 1];
#ifdef B200LSAP_TIMING
        long long tk = clock64();
#define NTICK(k)                                                                                                   \
    do                                                                                                             \
    {                                                                                                              \
        if (tid == 0)                                                                                              \
        {                                                                                                          \
            const long long now_ = clock64();                                                                      \
            acc##k += now_ - tk;                                                                                   \
            tk = now_;                                                                                             \
        }                                                                                                          \
    } while (0)
#else
#define NTICK(k)
#endif
#ifndef B200LSAP_NO_SOLO_ROUNDS
        if (n >= 2 && n <= kSoloMax)
        {
            // ---- two to four bidders: warp 0 alone plays the Jacobi rounds (bidders evaluated two at a time, the two
            // dependency chains interleaved; resolved in registers) for as long as at least two bidders remain: no block
            // barrier and no shared-memory exchange per round (a general round with <= 4 bidders cost 1900 cycles, two
            // barriers among them).  Same rounds as the general path: same winner rule, same order of the next list.
            if (warp == 0)
            {
                int r[kSoloMax];
#pragma unroll
                for (int u = 0; u < kSoloMax; u++)
                    r[u] = u < n ? q[u] : -1;
                int m = n, served = 0, served_bids = 0;
                for (;;)
                {
                    Top2<T> t[kSoloMax];
                    int prev[kSoloMax];
                    bool ok = true;
#pragma unroll
                    for (int u = 0; u < kSoloMax; u += 2)
                    {
                        if (u + 1 < m)
                        {
                            const int ii[2] = {r[u], r[u + 1]};
                            Top2<T> tt[2];
                            int pp[2];
                            bool hh[2];
                            narrow_eval2<kSmem, T>(s, prices_s, owner_s, ii, lane, tt, pp, hh);
                            t[u] = tt[0], t[u + 1] = tt[1];
                            prev[u] = pp[0], prev[u + 1] = pp[1];
                            ok = ok && hh[0] && hh[1];
                        }
                        else if (u < m)
                        {
                            ok = narrow_eval<kSmem, false, T>(s, prices_s, owner_s, r[u], lane, t[u], prev[u]) && ok;
                        }
                    }
                    if (!ok)
                        break; // a miss: the general round below scans the row
                    double bid[kSoloMax];
                    bool win[kSoloMax];
#pragma unroll
                    for (int u = 0; u < kSoloMax; u++)
                        bid[u] = u < m ? (double)t[u].a - t[u].w + eps : 0.0; // Auction.h:250
                    // highest bid per column, lowest row among equal bids
#pragma unroll
                    for (int u = 0; u < kSoloMax; u++)
                    {
                        win[u] = u < m;
#pragma unroll
                        for (int v = 0; v < kSoloMax; v++)
                            if (v != u && v < m && u < m && t[v].j == t[u].j &&
                                (bid[v] > bid[u] || (bid[v] == bid[u] && r[v] < r[u])))
                                win[u] = false;
                    }
                    if (lane == 0)
                    {
                        double top = 0.0;
#pragma unroll
                        for (int u = 0; u < kSoloMax; u++)
                            if (u < m)
                            {
                                s.profits[r[u]] = t[u].w - eps; // Auction.h:253
                                if (win[u])
                                {
                                    narrow_commit<kSmem, T>(s, prices_s, owner_s, r[u], t[u].j, bid[u], prev[u]);
                                    top = fmax(top, bid[u]);
                                }
                            }
                        if (top > sh.pmax)
                            sh.pmax = top;
                    }
                    served++;
                    served_bids += m;
                    // next list, in bidder order: a loser stays, a winner hands its place to the row it evicted
                    int nr[kSoloMax], mn = 0;
#pragma unroll
                    for (int u = 0; u < kSoloMax; u++)
                        nr[u] = -1;
#pragma unroll
                    for (int u = 0; u < kSoloMax; u++)
                    {
                        const int push = u < m ? (win[u] ? prev[u] : r[u]) : -1;
                        if (push >= 0)
                        {
#pragma unroll
                            for (int k = 0; k < kSoloMax; k++)
                                if (k == mn)
                                    nr[k] = push;
                            mn++;
                        }
                    }
#pragma unroll
                    for (int u = 0; u < kSoloMax; u++)
                        r[u] = nr[u];
                    m = mn;
                    __syncwarp(); // lane 0's shared-memory prices before the next gather
                    if (m < 2 || rounds + served >= budget || (esc_ok && work + served > work_budget))
                        break;
                }
                if (lane == 0)
                {
#pragma unroll
                    for (int u = 0; u < kSoloMax; u++)
                        if (u < m)
                            q[u] = r[u];
                    sh.qn = m;
                    sh.served = served;
                    sh.served_bids = served_bids;
                }
            }
            __syncthreads();
            const int served = sh.served;
            rounds += served;
            bids += sh.served_bids;
            work += served;
            solo_rounds += served;
            n = sh.qn;
            if (served > 0 || n < 2 || n > kSoloMax)
            {
                __syncthreads(); // sh.served is rewritten by the next episode
                continue;
            }
            // a miss in the very first of these rounds: fall through to the general round
        }
#endif
        if (n == 1)
        {
            // ---- one bidder: warp 0 alone, bid after bid along the eviction chain ------------------
            if (warp == 0)
            {
                int i = q[0], served = 0, left = 1;
#ifdef B200LSAP_TIMING
                long long evr[4] = {0, 0, 0, 0};
                long long* const evq = evr;
#else
                long long* const evq = nullptr;
#endif
                CacheEnt<T> e = s.cache[(long long)i * kCacheK + lane];
                RowCacheHdr h = s.cache_hdr[i];
                for (;;)
                {
                    Top2<T> t;
                    int prev;
                    if (!narrow_eval_regs<kSmem, true, T>(s, prices_s, owner_s, e, h, lane, t, prev, evq))
                        break; // the general round below scans the row
                    // the evicted row bids next: its cache is requested before this bid is committed,
                    // so the L2 latency of the load overlaps the stores
                    if (prev >= 0)
                    {
                        e = s.cache[(long long)prev * kCacheK + lane];
                        h = s.cache_hdr[prev];
                    }
                    if (lane == 0)
                    {
                        const double bid = (double)t.a - t.w + eps; // Auction.h:250
                        s.profits[i] = t.w - eps;                   // Auction.h:253
                        narrow_commit<kSmem, T>(s, prices_s, owner_s, i, t.j, bid, prev);
                        if (bid > sh.pmax)
                            sh.pmax = bid;
                    }
                    served++;
                    if (prev < 0)
                    {
                        left = 0;
                        break;
                    }
                    i = prev;
                    if (rounds + served >= budget || (esc_ok && work + served > work_budget))
                        break;
                    __syncwarp(); // lane 0's shared-memory prices before the next gather
                }
                if (lane == 0)
                {
                    q[0] = i;
                    sh.qn = left;
                    sh.served = served;
#ifdef B200LSAP_TIMING
                    for (int k = 0; k < 4; k++)
                        g_ev[k] += evr[k];
#endif
                }
            }
            NTICK(2);
            __syncthreads();
            const int served = sh.served;
            rounds += served;
            bids += served;
            fast += served;
            work += served;
            n = sh.qn;
            if (n == 0 || rounds >= budget || (esc_ok && work > work_budget))
            {
                __syncthreads(); // sh.served is rewritten by the next fast episode
                continue;
            }
            // a miss: fall through, the bidder q[0] takes the general round
        }
#ifdef B200LSAP_TIMING
        const long long t_round = clock64();
#endif
        // ---- bids from the caches: a warp per bidder -------------------------------------------------
        for (int b = warp; b < n; b += 2 * kNarrowWarps)
        {
            const int b1 = b + kNarrowWarps;
            Top2<T> t[2];
            int prev[2];
            bool hit[2];
            const int i[2] = {q[b], b1 < n ? q[b1] : q[b]};
            if (b1 < n)
                narrow_eval2<kSmem, T>(s, prices_s, owner_s, i, lane, t, prev, hit);
            else
            {
                // The row this bid evicts bids in the next round, and — a winner hands its place in the list to the row it
                // evicted — usually at this very position, i.e. in this very warp: its cache entry is requested now, so that
                // the L2 round trip runs behind this round's barriers and resolve instead of opening the next evaluation.
                CacheEnt<T> e;
                RowCacheHdr h;
                if (b == warp && pre_row == i[0])
                {
                    e = pre_e;
                    h = pre_h;
                }
                else
                {
                    e = s.cache[(long long)i[0] * kCacheK + lane];
                    h = s.cache_hdr[i[0]];
                }
                hit[0] = narrow_eval_regs<kSmem, false, T>(s, prices_s, owner_s, e, h, lane, t[0], prev[0], nullptr);
                if (b == warp)
                {
                    pre_row = prev[0];
                    if (prev[0] >= 0)
                    {
                        pre_e = s.cache[(long long)prev[0] * kCacheK + lane];
                        pre_h = s.cache_hdr[prev[0]];
                    }
                }
            }
            if (lane == 0)
            {
#pragma unroll
                for (int u = 0; u < 2; u++)
                {
                    const int bb = u ? b1 : b;
                    if (bb >= n)
                        break;
                    if (hit[u])
                    {
                        sh.col[bb] = t[u].j;
                        sh.prev[bb] = prev[u];
                        sh.bid[bb] = (double)t[u].a - t[u].w + eps; // Auction.h:250
                        s.profits[i[u]] = t[u].w - eps;             // Auction.h:253
                    }
                    else
                    {
                        sh.col[bb] = kMissMark;
                        atomicAdd(&sh.nmiss, 1);
                    }
                }
            }
        }
        NTICK(3);
        __syncthreads();
        NTICK(4);
        const int nmiss = sh.nmiss;
        if (nmiss > 0)
        {
            if (kSmem)
                narrow_flush_prices(s, prices_s, owner_s); // the row scans read prices from global memory
            narrow_misses<kSmem, T>(s, sh, owner_s, q, n, eps, amax); // ends with a barrier
            if (tid == 0)
                sh.nmiss = 0;
            if (sh.stop)
                break;
            misses += nmiss;
            work += 4 * nmiss;
        }
        work++;
        // ---- resolve: highest bid per column, lowest row among equal bids ------------------------------
        if (n <= 32)
        {
            // one warp, in registers: match.any groups the bidders by column; only contested columns
            // (rare) cost a reduction each
            if (warp == 0)
            {
                const bool act = lane < n;
                const int my_row = act ? q[lane] : -1;
                const int my_col = act ? sh.col[lane] : -1 - lane;
                const double my_bid = act ? sh.bid[lane] : 0.0;
                int my_prev = act ? sh.prev[lane] : -1;
                // columns first (match.any costs ~11 cycles per distinct value on sm_100: 370 for 32
                // distinct columns); only a round with a contested column compares bids
                bool dup = false;
                for (int k0 = 0; k0 < n; k0 += 4) // four independent shuffles in flight
                {
#pragma unroll
                    for (int u = 0; u < 4; u++)
                    {
                        const int k = k0 + u;
                        const int oc = __shfl_sync(kFullMask, my_col, k & 31);
                        dup |= k < n && oc == my_col && k != lane;
                    }
                }
                bool win = act;
                if (__any_sync(kFullMask, dup))
                    for (int k = 0; k < n; k++)
                    {
                        const int oc = __shfl_sync(kFullMask, my_col, k);
                        const double ob = __shfl_sync(kFullMask, my_bid, k);
                        const int orow = __shfl_sync(kFullMask, my_row, k);
                        if (k != lane && oc == my_col && (ob > my_bid || (ob == my_bid && orow < my_row)))
                            win = false;
                    }
                NTICK(7);
                if (win)
                {
                    narrow_commit<kSmem, T>(s, prices_s, owner_s, my_row, my_col, my_bid, my_prev);
                    if (my_bid > sh.pmax)
                        atomicMax(reinterpret_cast<unsigned long long*>(&sh.pmax),
                                  (unsigned long long)__double_as_longlong(my_bid));
                }
                // next list, in bidder order: a loser stays, a winner hands its place to the row it evicted
                const int push = act ? (win ? my_prev : my_row) : -1;
                const unsigned mask = __ballot_sync(kFullMask, push >= 0);
                if (push >= 0)
                    qn[__popc(mask & ((1u << lane) - 1))] = push;
                if (lane == 0)
                    sh.qn = __popc(mask);
            }
        }
        else
        {
            // per column the best bid through a hash table keyed by column: insert, then a compare-and-swap
            // argmax on the slot (highest bid, lowest row among equal bids); O(1) per bidder instead of a
            // pass over all bidders
            int my_row = -1, my_col = -1, my_prev = -1, my_h = -1;
            double my_bid = 0.0;
            bool win = false;
            if (tid < n)
            {
                my_row = q[tid];
                my_col = sh.col[tid];
                my_bid = sh.bid[tid];
                int h = (int)(((unsigned)my_col * 2654435761u) >> (32 - kNarrowHashBits));
                for (;;)
                {
                    const int old = atomicCAS(&sh.hcol[h], -1, my_col);
                    if (old == -1 || old == my_col)
                        break;
                    h = (h + 1) & (kNarrowHash - 1);
                }
                my_h = h;
                int cur = *(volatile int*)&sh.hbest[h];
                for (;;)
                {
                    if (cur != -1)
                    {
                        const double ob = sh.bid[cur];
                        if (!(my_bid > ob || (my_bid == ob && my_row < q[cur])))
                            break;
                    }
                    const int old = atomicCAS(&sh.hbest[h], cur, tid);
                    if (old == cur)
                        break;
                    cur = old;
                }
            }
            __syncthreads(); // every bid has met its column's slot
            if (tid < n)
            {
                win = sh.hbest[my_h] == tid;
                if (win)
                {
                    my_prev = sh.prev[tid];
                    narrow_commit<kSmem, T>(s, prices_s, owner_s, my_row, my_col, my_bid, my_prev);
                    if (my_bid > sh.pmax)
                        atomicMax(reinterpret_cast<unsigned long long*>(&sh.pmax),
                                  (unsigned long long)__double_as_longlong(my_bid));
                }
            }
            const int push = tid < n ? (win ? my_prev : my_row) : -1;
            const unsigned mask = __ballot_sync(kFullMask, push >= 0);
            if (lane == 0)
                sh.warp_count[warp] = __popc(mask);
            __syncthreads(); // warp_count is complete, every winner has been read from the table
            if (my_h >= 0)
            {
                sh.hcol[my_h] = -1;
                sh.hbest[my_h] = -1;
            }
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
        rounds++;
        bids += n;
        buf ^= 1;
        NTICK(5);
        __syncthreads();
        NTICK(6);
#ifdef B200LSAP_TIMING
        {
            // whole general round by number of bidders: 1-4, 5-16, 17-32, 33-256
            const long long dt = clock64() - t_round;
            if (n <= 4) { bk_c0 += dt; bk_r0++; }
            else if (n <= 16) { bk_c1 += dt; bk_r1++; }
            else if (n <= 32) { bk_c2 += dt; bk_r2++; }
            else { bk_c3 += dt; bk_r3++; }
        }
#endif
    }
    __syncthreads();
    if (kSmem)
        narrow_writeback(s, prices_s, owner_s);
    if (tid == 0)
    {
        const bool finished = sh.qn == 0 && !sh.stop && ctrl->status == 0;
        if (finished)
        {
            ctrl->count[cur] = 0;
            if (ctrl->phase + 1 >= ctrl->n_phases)
                ctrl->done = 1;
        }
        else if (escalated && ctrl->status == 0)
            escalate_schedule(ctrl, cur);
        else
            ctrl->phase_work = work;
        ctrl->rounds += rounds;
        ctrl->bids += bids;
        ctrl->narrow_rounds += rounds;
        ctrl->cache_hits += bids - misses;
        ctrl->cache_misses += misses;
        ctrl->cache_fresh = misses * 64 <= bids ? 1 : 0; // next phase: keep the caches unless this one missed > 1.5 %
        ctrl->gs_bids += fast; // bids placed by the one-bidder path (included in rounds and bids)
        ctrl->dbg[12] += solo_rounds; // rounds with 2..kSoloMax bidders played by warp 0 alone (b200lsap_debug_counters)
        if (sh.pmax > ctrl->pmax)
            ctrl->pmax = sh.pmax;
#ifdef B200LSAP_TIMING
        ctrl->dbg[0] += clock64() - t_begin;
        ctrl->dbg[8] += rounds;
        ctrl->dbg[9] += bids;
        ctrl->dbg[1] += fast;
#if B200LSAP_VAR == 13 // sections of a one-bidder bid: wait for the cache entry, price gather, reductions, winner's data
        for (int k = 0; k < 4; k++)
        {
            ctrl->dbg[2 + k] += g_ev[k];
            g_ev[k] = 0;
        }
        ctrl->dbg[6] += acc2;
#elif B200LSAP_VAR == 12 // rounds and cycles of the general rounds by number of bidders
        ctrl->dbg[2] += bk_r0; ctrl->dbg[3] += bk_c0;
        ctrl->dbg[4] += bk_r1; ctrl->dbg[5] += bk_c1;
        ctrl->dbg[6] += bk_r2; ctrl->dbg[7] += bk_c2;
        ctrl->dbg[10] += bk_r3; ctrl->dbg[11] += bk_c3;
#else
        ctrl->dbg[2] += acc2;
        ctrl->dbg[3] += acc3;
        ctrl->dbg[4] += acc4;
        ctrl->dbg[5] += acc5;
        ctrl->dbg[6] += acc6;
        ctrl->dbg[7] += acc7;
        ctrl->dbg[10] += bk_r0; ctrl->dbg[11] += bk_c0;
#endif
#endif
    }
}

} // namespace b200lsap
