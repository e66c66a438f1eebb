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
// will read) CTA 0 finishes the phase alone with block barriers while the other CTAs wait at one
// grid sync.
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
    return (int)min((long long)kThreads, max(32ll, (256ll * 1024) / avg));
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
    __shared__ int sh_n;

    int cur = 0, phase = 0;
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
            if (blockIdx.x == 0)
            {
                int nn = n, c = cur;
                long long r = 0, b = 0;
                while (nn > 0 && spent + r < max_rounds && ctrl->status == 0)
                {
                    if (threadIdx.x == 0)
                        ctrl->count[c ^ 1] = 0;
                    const double pmax = fmax(ctrl->pmax, amax);
                    for (int k = threadIdx.x >> 5; k < nn; k += kWarpsPerBlock)
                        csr_bid_row<T>(s, m, s.list[c][k], eps, pmax, lane);
                    __syncthreads();
                    if (ctrl->status != 0)
                        break;
                    resolve_phase<T>(s, s.list[c], nn, s.list[c ^ 1], &s.ctrl->count[c ^ 1], threadIdx.x, kThreads, lane);
                    __syncthreads();
                    if (threadIdx.x == 0)
                        sh_n = ctrl->count[c ^ 1];
                    __syncthreads();
                    r++;
                    b += nn;
                    nn = sh_n;
                    c ^= 1;
                }
                if (threadIdx.x == 0)
                {
                    ctrl->cur = c;
                    ctrl->narrow_rounds += r;
                    ctrl->gs_bids += b; // bids of the narrow episodes (added to bids at the end)
                    if (nn > 0 && ctrl->status == 0)
                        ctrl->status = 3;
                    __threadfence();
                }
            }
            grid.sync();
            if (ctrl->status != 0)
                break;
            cur = ctrl->cur;
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
