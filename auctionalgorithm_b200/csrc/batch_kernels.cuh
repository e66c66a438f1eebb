// batch_kernels.cuh — batched mode for MHT-style data association (north_star (4)):
// many independent small problems (rows <= 32, cols <= 64), one WARP per problem, the matrix
// of the problem held in shared memory, prices / owners / assignment held in registers.
//
// Unlike the big dense solver this kernel keeps the reference's own schedule: it is the
// Gauss-Seidel sweep of Auction::forward (Auction.h:210-306) — rows bid one after the other in
// increasing row order, each seeing the prices left by the previous bid, sweeps repeat until
// no row is unassigned (Auction.h:66-68) — so for a given problem it places exactly the bids
// the reference places and ends with exactly the reference's assignment, ties included.
// What is parallel is the inside of findForward (Auction.h:124-152): lane l owns columns l
// and l+32, the row scan is one shared-memory read per lane and the top-2 is a warp reduction.
#pragma once

#include "common.cuh"

namespace b200lsap
{

constexpr int kBatchWarps = 8; // warps (= problems in flight) per CTA
constexpr int kBatchMaxRows = 32;
constexpr int kBatchMaxCols = 64;

struct BatchParams
{
    const float* a;   // [batch][rows*cols], row-major inside a problem
    int* row2col;     // [batch][rows]
    long long batch;
    int rows, cols;
    double eps;       // eps_multiplier / cols (Auction.h:55)
    int max_bids;     // watchdog per problem
    int* status;      // device flag, set to a b200lsap_status on failure
};

__global__ void __launch_bounds__(kBatchWarps * 32) k_auction_batch(BatchParams P)
{
    extern __shared__ __align__(16) float smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int n = P.rows * P.cols;
    const int n_pad = (n + 3) & ~3;
    float* sA = smem + (size_t)warp * n_pad;
    const long long warp_global = (long long)blockIdx.x * kBatchWarps + warp;
    const long long total_warps = (long long)gridDim.x * kBatchWarps;
    const bool has0 = lane < P.cols, has1 = lane + 32 < P.cols;
    const unsigned all_rows = P.rows == 32 ? kFullMask : ((1u << P.rows) - 1u);

    for (long long prob = warp_global; prob < P.batch; prob += total_warps)
    {
        const float* g = P.a + prob * n;
        if ((n & 3) == 0)
        {
            for (int k = lane; k < n / 4; k += 32)
                reinterpret_cast<float4*>(sA)[k] = ld_stream4(g + 4 * k);
        }
        else
        {
            for (int k = lane; k < n; k += 32)
                sA[k] = __ldg(g + k);
        }
        __syncwarp();

        double p0 = 0.0, p1 = 0.0; // prices of columns lane, lane+32 (Auction.h:58)
        int o0 = -1, o1 = -1;      // rows holding them
        int mycol = -1;            // lane as a row: its column
        unsigned unassigned = all_rows; // warp-uniform
        int bids = 0;
        bool infeasible = false;

        while (unassigned)
        {
            unsigned todo = unassigned; // one forward sweep (Auction.h:220): rows in increasing order
            while (todo)
            {
                const int i = __ffs(todo) - 1;
                todo &= todo - 1;
                // findForward (Auction.h:124-152), two columns per lane
                Top2<float> t = top2_empty<float>();
                if (has0)
                    top2_push(t, sA[i * P.cols + lane], p0, lane);
                if (has1)
                    top2_push(t, sA[i * P.cols + 32 + lane], p1, lane + 32);
                top2_warp_reduce(t);
                if (t.j < 0)
                {
                    infeasible = true;
                    unassigned = 0;
                    break;
                }
                const double bid = (double)t.a - t.w + P.eps; // Auction.h:250
                // lambda = 0 <= bid always holds here (Auction.h:257): assign, evict the owner
                int prev = -1;
                if (lane == (t.j & 31))
                {
                    if (t.j < 32)
                    {
                        prev = o0;
                        p0 = bid;
                        o0 = i;
                    }
                    else
                    {
                        prev = o1;
                        p1 = bid;
                        o1 = i;
                    }
                }
                prev = __shfl_sync(kFullMask, prev, t.j & 31);
                if (lane == i)
                    mycol = t.j;
                unassigned &= ~(1u << i);
                if (prev >= 0)
                {
                    if (lane == prev)
                        mycol = -1;
                    unassigned |= 1u << prev;
                    if (prev > i)
                        todo |= 1u << prev; // unlocked ahead of the sweep position: visited in this sweep
                }
                if (++bids > P.max_bids)
                {
                    infeasible = true;
                    unassigned = 0;
                    break;
                }
            }
        }
        if (infeasible && lane == 0)
            *P.status = bids > P.max_bids ? 3 : 2;
        if (lane < P.rows)
            P.row2col[prob * P.rows + lane] = mycol;
        __syncwarp();
    }
}

// Column-major problems (Eigen's default) -> row-major problems, one pass over the batch.
__global__ void __launch_bounds__(256) k_batch_to_row_major(const float* __restrict__ in, float* __restrict__ out,
                                                            long long batch, int rows, int cols)
{
    const long long n = (long long)rows * cols;
    const long long total = batch * n;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x)
    {
        const long long b = e / n;
        const int r = (int)(e - b * n);
        const int i = r / cols, j = r - i * cols;
        out[e] = in[b * n + (long long)j * rows + i];
    }
}

} // namespace b200lsap
