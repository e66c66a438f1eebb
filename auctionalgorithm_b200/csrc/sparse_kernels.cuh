// sparse_kernels.cuh — the forward auction on a compressed row-major sparse matrix (CSR).
//
// Included by capi.cu after dense_kernels.cuh (uses DenseState's price / owner / slot / list arrays,
// submit_bid, resolve_phase and reset_phase: everything of a round except the row scan is the dense
// path's, so rounds, tie rules and prices are those of a dense solve of the same matrix with its
// missing entries as "no edge").
//
// Replaces Auction<Scalar, Eigen::SparseMatrix<Scalar, Eigen::RowMajor>>::solve: findForward on the
// stored entries of a row (Auction.h:166-198) under the controller of Auction.h:44-109.  Every stored
// entry is an edge (the reference skips nothing here, unlike its dense scan); columns are ascending
// inside a row, as in Eigen's compressed format, so "first in storage order" = "lowest column" for
// equal values.
//
// One cooperative launch per solve.  A round with many bidders is a warp per bidding row over the
// whole grid (lanes stride the row's entries: 8- or 12-byte (column, value) pairs, coalesced; prices
// are gathered from L2) with two grid syncs.  Rounds only ever get narrower inside an eps phase, so
// once a round's bidders fit one CTA (csr_narrow_limit: by their number and by the entries they
// will read) CTA 0 finishes the phase alone (csr_narrow_episode) while the other CTAs wait at one
// grid sync: bidder list, bids and the per-column resolve live in shared memory (a hash table keyed
// by column with a compare-and-swap argmax: highest bid, lowest row among equal bids), the matrix
// entries come through L1 (read-only for the whole solve), prices and owners are only written by
// this CTA, so a round is three block barriers and no global atomic.
#pragma once

namespace b200lsap
{

template <typename T> struct CsrView
{
    const int* ptr; // [rows + 1]
    const int* col; // [nnz], ascending inside a row
    const T* val;   // [nnz]
    long long nnz;
};

// max |value| -> ctrl->amax (ordered as u64 bits: values are >= 0 after fabs)
template <typename T> __global__ void __launch_bounds__(256) k_absmax_csr(const T* __restrict__ val, long long nnz, DenseCtrl* ctrl)
{
    T m = T(0);
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < nnz; e += (long long)gridDim.x * blockDim.x)
        m = fmax(m, fabs(val[e]));
    double d = (double)m;
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1)
        d = fmax(d, __shfl_xor_sync(kFullMask, d, s));
    if ((threadIdx.x & 31) == 0 && d > 0.0)
        atomicMax(reinterpret_cast<unsigned long long*>(&ctrl->amax), (unsigned long long)__double_as_longlong(d));
}

// Auction.h:174-195 for one lane's share of the row; stored zeros are edges
template <typename T> __device__ __forceinline__ void top2_push_stored(Top2<T>& t, T a, double p, int col)
{
    double d;
    if constexpr (sizeof(T) == 4)
        d = widen_exact(a) - p;
    else
        d = a - p;
    const bool gv = d > t.v;
    const bool gw = d > t.w;
    t.w = gv ? t.v : (gw ? d : t.w);
    t.v = gv ? d : t.v;
    t.j = gv ? col : t.j;
    t.a = gv ? a : t.a;
}

// One warp: best / second best of row i over its stored entries, every lane ends with the result.
template <typename T, bool kDeep>
__device__ __forceinline__ Top2<T> csr_scan_row(const CsrView<T>& m, const double* prices, int i, int lane)
{
    Top2<T> t = top2_empty<T, kDeep>();
    const int k0 = m.ptr[i], k1 = m.ptr[i + 1];
    for (int k = k0 + lane; k < k1; k += 64) // two independent gathers in flight per lane
    {
        const int j0 = __ldg(m.col + k);
        const T a0 = __ldg(m.val + k);
        const bool two = k + 32 < k1;
        const int j1 = two ? __ldg(m.col + k + 32) : 0;
        const T a1 = two ? __ldg(m.val + k + 32) : T(0);
        const double p0 = __ldcg(prices + j0);
        const double p1 = two ? __ldcg(prices + j1) : 0.0;
        top2_push_stored(t, a0, p0, j0);
        if (two)
            top2_push_stored(t, a1, p1, j1);
    }
    top2_warp_reduce(t);
    return t;
}

// All lanes: the bid of row i (the -1e6 rescue of the dense path's rescue_scan included)
template <typename T>
__device__ __forceinline__ void csr_bid_row(const DenseState<T>& s, const CsrView<T>& m, int i, double eps, double pmax,
                                            int lane)
{
    Top2<T> t = csr_scan_row<T, false>(m, s.prices, i, lane);
    if (t.j < 0 && m.ptr[i + 1] > m.ptr[i])
    {
        t = csr_scan_row<T, true>(m, s.prices, i, lane);
        if (t.j >= 0 && t.w == -CUDART_INF)
            t.w = fmin(-kNoSecondBest, t.v);
    }
    if (lane == 0)
        submit_bid(s, i, t, eps, pmax);
}

// Bidders that one CTA finishes faster than the grid would: at most its threads (one per bidder in
// the resolve), fewer when rows are long.
__device__ __forceinline__ int csr_narrow_limit(long long nnz, int rows)
{
    const long long avg = max(1ll, nnz / max(rows, 1));
    return (int)min((long long)kThreads, max(1ll, (64ll * 1024) / avg));
}

// ---------------------------------------------------------------------------------------------
// One CTA finishes an eps phase (n <= kThreads bidders; rounds never get wider inside a phase).
// ---------------------------------------------------------------------------------------------
constexpr int kCsrHash = 2048; // >= 2 kThreads slots: open addressing stays short

template <typename T> struct CsrNarrowShared
{
    int q[2][kThreads];      // bidder rows of this round / the next
    int bcol[kThreads];      // column bidder k bids for
    double bbid[kThreads];   // its bid
    int bown[kThreads];      // owner of that column when the bid was made (unchanged until the commit)
    int hcol[kCsrHash];      // hash table: column, -1 empty
    int hbest[kCsrHash];     //             bidder index of the best bid so far, -1 none
    Top2<T> part[kWarpsPerBlock]; // long rows: partial results of the row's segments
    int qn[2];               // bidders in q[0] / q[1]
    int served;              // bids placed by the one-bidder chain of warp 0
    int status;
    double pmax;
};

// bidder x is a better bid than bidder y for the same column (Jacobi resolve: highest bid, lowest row)
template <typename T> __device__ __forceinline__ bool csr_better(const CsrNarrowShared<T>& sh, const int* q, int x, int y)
{
    const double bx = sh.bbid[x], by = sh.bbid[y];
    return bx > by || (bx == by && q[x] < q[y]);
}

// Warp top-2 of per-lane results on the ordered bits of the doubles (narrow_kernel.cuh): five
// redux.sync instead of five levels of 64-bit shuffles and dependent fp64 compares.  Same result as
// top2_warp_reduce: the best is the largest v, lowest column among equals; the second best is the
// largest of the other lanes' v and the winner's own w.  Lanes without an entry hold the start
// value (-1e6 or -inf) and column -1.
template <typename T> __device__ __forceinline__ void top2_warp_reduce_keys(Top2<T>& t)
{
    const unsigned long long kv = ordered_key(t.v);
    const unsigned hi = (unsigned)(kv >> 32), lo = (unsigned)kv;
    const bool has = t.j >= 0;
    const unsigned mh = __reduce_max_sync(kFullMask, has ? hi : 0u);
    const bool c1 = has && hi == mh;
    const unsigned ml = __reduce_max_sync(kFullMask, c1 ? lo : 0u);
    const bool top = c1 && lo == ml;
    const int mj = (int)__reduce_min_sync(kFullMask, top ? (unsigned)t.j : 0xffffffffu);
    const bool winner = top && t.j == mj;
    const unsigned long long k2 = ordered_key(winner ? t.w : t.v); // lanes without an entry: the start value
    const unsigned hi2 = (unsigned)(k2 >> 32), lo2 = (unsigned)k2;
    const unsigned mh2 = __reduce_max_sync(kFullMask, hi2);
    const unsigned ml2 = __reduce_max_sync(kFullMask, hi2 == mh2 ? lo2 : 0u);
    const unsigned wmask = __ballot_sync(kFullMask, winner);
    if (wmask == 0)
        return; // no lane holds an entry: every lane keeps its (identical) empty state
    const int src = __ffs(wmask) - 1;
    t.a = __shfl_sync(kFullMask, t.a, src);
    t.j = mj;
    t.v = key_to_double(((unsigned long long)mh << 32) | ml);
    t.w = key_to_double(((unsigned long long)mh2 << 32) | ml2);
}

// One warp: top-2 of entries [k0, k1) of a row (a segment or all of it), every lane ends with it.
template <typename T, bool kDeep>
__device__ __forceinline__ Top2<T> csr_scan_range(const CsrView<T>& m, const double* prices, int k0, int k1, int lane)
{
    Top2<T> t = top2_empty<T, kDeep>();
    for (int k = k0 + lane; k < k1; k += 64) // two independent gathers in flight per lane
    {
        const int j0 = __ldg(m.col + k);
        const T a0 = __ldg(m.val + k);
        const bool two = k + 32 < k1;
        const int j1 = two ? __ldg(m.col + k + 32) : 0;
        const T a1 = two ? __ldg(m.val + k + 32) : T(0);
        // prices are written by this CTA only during the episode: plain loads (L1) stay coherent
        const double p0 = prices[j0];
        const double p1 = two ? prices[j1] : 0.0;
        top2_push_stored(t, a0, p0, j0);
        if (two)
            top2_push_stored(t, a1, p1, j1);
    }
    top2_warp_reduce_keys(t);
    return t;
}

// All lanes of one warp hold the row's top-2: place bidder k's bid in shared memory.
template <typename T>
__device__ __forceinline__ void csr_post_bid(const DenseState<T>& s, const CsrView<T>& m, CsrNarrowShared<T>& sh, int k, int i,
                                             Top2<T> t, double eps, int lane)
{
    if (t.j < 0 && m.ptr[i + 1] > m.ptr[i])
    {
        // every value at or below -1e6: the rescue of the dense path's rescue_scan
        t = csr_scan_range<T, true>(m, s.prices, m.ptr[i], m.ptr[i + 1], lane);
        if (t.j >= 0 && t.w == -CUDART_INF)
            t.w = fmin(-kNoSecondBest, t.v);
    }
    if (lane == 0)
    {
        if (t.j < 0)
        {
            sh.status = 2; // B200LSAP_ERR_INFEASIBLE
            sh.bcol[k] = -1;
            return;
        }
        const double bid = (double)t.a - t.w + eps; // Auction.h:250
        s.profits[i] = t.w - eps;                   // Auction.h:253
        sh.bcol[k] = t.j;
        sh.bbid[k] = bid;
        sh.bown[k] = s.owner[t.j]; // the load runs beside the barrier and the resolve
        if (bid > sh.pmax)
            atomicMax(reinterpret_cast<unsigned long long*>(&sh.pmax), (unsigned long long)__double_as_longlong(bid));
    }
}

// Returns the rounds played; *bids_out their bidders.  On return sh.qn[buf] bidders are left (0: phase done).
template <typename T>
__device__ __noinline__ long long csr_narrow_episode(const DenseState<T>& s, const CsrView<T>& m, CsrNarrowShared<T>& sh,
                                                     const int* list, int n0, double eps, double pmax0, long long budget,
                                                     long long* bids_out)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int k = tid; k < kCsrHash; k += kThreads)
    {
        sh.hcol[k] = -1;
        sh.hbest[k] = -1;
    }
    if (tid < n0)
        sh.q[0][tid] = list[tid];
    if (tid == 0)
    {
        sh.qn[0] = n0;
        sh.qn[1] = 0;
        sh.status = 0;
        sh.pmax = pmax0;
    }
    __syncthreads();
    const long long avg = max(1ll, m.nnz / max(s.rows, 1));
    int buf = 0, my_h = -1;
    long long rounds = 0, bids = 0;
    for (;;)
    {
        const int n = sh.qn[buf];
        if (n == 0 || rounds >= budget)
            break;
        if (n == 1)
        {
            // ---- one bidder: warp 0 alone, bid after bid along the eviction chain (the dense path's k_narrow does the
            // same): scan the row, bid, evict, go on with the evicted row; no block barrier, no hash table.  These are
            // the same Jacobi rounds (a round with one bidder has one winner), so prices come out identical.
            if (warp == 0)
            {
                int i = sh.q[buf][0], served = 0, left = 1;
                for (;;)
                {
                    const Top2<T> t = csr_scan_range<T, false>(m, s.prices, m.ptr[i], m.ptr[i + 1], lane);
                    if (t.j < 0)
                        break; // every value at or below -1e6, or no entry: the general round below deals with it
                    const double bid = (double)t.a - t.w + eps; // Auction.h:250
                    const int prev = s.owner[t.j];
                    __syncwarp();
                    if (lane == 0)
                    {
                        s.profits[i] = t.w - eps; // Auction.h:253
                        s.owner[t.j] = i;
                        s.prices[t.j] = bid;      // Auction.h:259
                        s.row2col[i] = t.j;
                        if (prev >= 0)
                            s.row2col[prev] = -1; // Auction.h:270
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
                    if (rounds + served >= budget)
                        break;
                    __syncwarp(); // lane 0's price and owner before the next row's gathers
                }
                if (lane == 0)
                {
                    sh.q[buf][0] = i;
                    sh.qn[buf] = left;
                    sh.served = served;
                }
            }
            __syncthreads();
            const int served = sh.served;
            rounds += served;
            bids += served;
            if (served > 0)
            {
                __syncthreads(); // sh.served is rewritten by the next chain
                continue;
            }
            // nothing placed (a row that needs the rescue scan): the general round below takes it
        }
        const int* q = sh.q[buf];
        int* qnext = sh.q[buf ^ 1];
        // ---- bids ----------------------------------------------------------------------------------
        // long rows and few bidders: S warps share a row (segments of >= 64 entries on average)
        int S = 1;
        while (S * 2 * n <= kWarpsPerBlock && avg >= 128ll * S)
            S *= 2;
        if (S == 1)
        {
            for (int k = warp; k < n; k += kWarpsPerBlock)
            {
                const int i = q[k];
                const Top2<T> t = csr_scan_range<T, false>(m, s.prices, m.ptr[i], m.ptr[i + 1], lane);
                csr_post_bid<T>(s, m, sh, k, i, t, eps, lane);
            }
        }
        else
        {
            const int k = warp / S, sg = warp - k * S;
            if (k < n)
            {
                const int i = q[k];
                const int r0 = m.ptr[i], r1 = m.ptr[i + 1];
                const int len = (r1 - r0 + S - 1) / S;
                const int k0 = min(r0 + sg * len, r1), k1 = min(k0 + len, r1);
                const Top2<T> t = csr_scan_range<T, false>(m, s.prices, k0, k1, lane);
                if (lane == 0)
                    sh.part[warp] = t;
            }
            __syncthreads();
            if (warp < n)
            {
                Top2<T> t = lane < S ? sh.part[warp * S + lane] : top2_empty<T>();
                top2_warp_reduce(t);
                csr_post_bid<T>(s, m, sh, warp, q[warp], t, eps, lane);
            }
        }
        if (tid == 0)
            sh.qn[buf ^ 1] = 0;
        __syncthreads();
        if (sh.status != 0)
            break;
        // ---- resolve: one thread per bidder; per column the best bid by compare-and-swap ----------------
        int col = -1;
        if (tid < n)
        {
            col = sh.bcol[tid];
            int h = (int)(((unsigned)col * 2654435761u) >> 21); // top 11 bits: kCsrHash slots
            for (;;)
            {
                const int old = atomicCAS(&sh.hcol[h], -1, col);
                if (old == -1 || old == col)
                    break;
                h = (h + 1) & (kCsrHash - 1);
            }
            my_h = h;
            int cur = *(volatile int*)&sh.hbest[h];
            for (;;)
            {
                if (cur != -1 && !csr_better(sh, q, tid, cur))
                    break;
                const int old = atomicCAS(&sh.hbest[h], cur, tid);
                if (old == cur)
                    break;
                cur = old;
            }
        }
        __syncthreads();
        int push = -1;
        if (tid < n)
        {
            const int row = q[tid];
            if (sh.hbest[my_h] == tid)
            {
                const int prev = sh.bown[tid];
                s.owner[col] = row;
                s.prices[col] = sh.bbid[tid]; // Auction.h:259
                s.row2col[row] = col;
                if (prev >= 0)
                    s.row2col[prev] = -1;     // Auction.h:270
                push = prev;
            }
            else
                push = row;
        }
        const unsigned mask = __ballot_sync(kFullMask, push >= 0);
        if (mask)
        {
            int base = 0;
            if (lane == 0)
                base = atomicAdd(&sh.qn[buf ^ 1], __popc(mask));
            base = __shfl_sync(kFullMask, base, 0);
            if (push >= 0)
                qnext[base + __popc(mask & ((1u << lane) - 1))] = push;
        }
        __syncthreads();
        if (my_h >= 0)
        {
            sh.hcol[my_h] = -1; // next round's inserts come after its first barrier
            sh.hbest[my_h] = -1;
            my_h = -1;
        }
        rounds++;
        bids += n;
        buf ^= 1;
    }
    __syncthreads();
    if (tid == 0)
    {
        sh.qn[0] = sh.qn[buf]; // what is left, for the caller
        if (sh.pmax > s.ctrl->pmax)
            s.ctrl->pmax = sh.pmax;
        if (sh.status != 0)
            s.ctrl->status = sh.status;
    }
    __syncthreads();
    *bids_out = bids;
    return rounds;
}

template <typename T>
__global__ void __launch_bounds__(kThreads, 1) k_auction_csr(const __grid_constant__ DenseState<T> s,
                                                             const __grid_constant__ CsrView<T> m)
{
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31;
    const int warp_global = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const int total_warps = gridDim.x * kWarpsPerBlock;
    const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
    const long long n_threads = (long long)gridDim.x * kThreads;
    volatile DenseCtrl* ctrl = s.ctrl;
    __shared__ CsrNarrowShared<T> nsh;

    int cur = 0, phase = 0;
    int solo = ctrl->solo_seq; // narrow episodes so far (the same in every CTA)
    const int n_phases = ctrl->n_phases;
    const long long max_rounds = ctrl->max_rounds;
    long long wide_rounds = 0, wide_bids = 0;
    double eps = ctrl->eps[0];
    const double amax = ctrl->amax;
    const int narrow_limit = csr_narrow_limit(m.nnz, s.rows);

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
        // narrow_rounds / gs_bids are CTA 0's, published before the grid sync that ends a narrow episode
        const long long spent = wide_rounds + ctrl->narrow_rounds;
        if (spent >= max_rounds)
        {
            if (tid == 0)
                ctrl->status = 3; // B200LSAP_ERR_MAX_ROUNDS
            break;
        }
        if (n <= narrow_limit)
        {
            // ---- CTA 0 finishes the phase: the same rounds with block barriers ----------------------
            solo++;
            if (blockIdx.x != 0)
                solo_wait(s.ctrl, solo); // doze instead of hammering the grid barrier (dense_kernels.cuh)
            if (blockIdx.x == 0)
            {
                long long b = 0;
                const long long r = csr_narrow_episode<T>(s, m, nsh, s.list[cur], n, eps, fmax(ctrl->pmax, amax),
                                                          max_rounds - spent, &b);
                if (threadIdx.x == 0)
                {
                    ctrl->narrow_rounds += r;
                    ctrl->gs_bids += b; // bids of the narrow episodes (added to bids at the end)
                    if (nsh.qn[0] > 0 && ctrl->status == 0)
                        ctrl->status = 3; // round budget spent
                    ctrl->count[cur] = 0;
                    __threadfence();
                }
                solo_done(s.ctrl, solo);
            }
            grid.sync();
            if (ctrl->status != 0)
                break;
            continue;
        }
        // ---- wide round: a warp per bidding row over the whole grid ---------------------------------
        if (tid == 0)
            ctrl->count[cur ^ 1] = 0;
        {
            const double pmax = fmax(ctrl->pmax, amax);
            const int* __restrict__ list = s.list[cur];
            for (int b = warp_global; b < n; b += total_warps)
                csr_bid_row<T>(s, m, list[b], eps, pmax, lane);
        }
        grid.sync();
        if (ctrl->status != 0) // raised only inside a bid phase: every thread reads the same value here
            break;
        resolve_phase<T>(s, s.list[cur], n, s.list[cur ^ 1], &s.ctrl->count[cur ^ 1], tid, n_threads, lane);
        grid.sync();
        cur ^= 1;
        wide_rounds++;
        wide_bids += n;
    }
    if (tid == 0)
    {
        ctrl->rounds += wide_rounds + ctrl->narrow_rounds;
        ctrl->bids += wide_bids + ctrl->gs_bids;
        ctrl->gs_bids = 0;
        ctrl->phase = phase;
        ctrl->cur = cur;
        ctrl->done = 1;
    }
}

} // namespace b200lsap
