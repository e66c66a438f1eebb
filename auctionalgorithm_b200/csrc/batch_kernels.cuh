// batch_kernels.cuh — batched mode for MHT-style data association (north_star (4)):
// many independent small problems (rows <= 32, cols <= 64), the matrix of a problem held in
// shared memory, prices / owners / assignment held in registers.
//
// Unlike the big dense solver this kernel keeps the reference's own schedule: it is the
// Gauss-Seidel sweep of Auction::forward (Auction.h:210-306) — rows bid one after the other in
// increasing row order, each seeing the prices left by the previous bid, sweeps repeat until
// no row is unassigned (Auction.h:66-68) — so for a given problem it places exactly the bids
// the reference places and ends with exactly the reference's assignment, ties included.
// What is parallel is the inside of findForward (Auction.h:124-152) and the problems themselves:
// a warp solves FOUR problems at once, eight lanes per problem, lane l of a group owning columns
// l, l+8, ..., l+56.  The row scan is then up to eight shared-memory reads per lane and the top-2
// a three-stage shuffle reduction inside the group; one instruction stream serves four bids.
// The four groups stay converged: every iteration of the loop is one bid of every group that
// still has an unassigned row (a finished group idles under a predicate).
#pragma once

#include "common.cuh"

namespace b200lsap
{

constexpr int kBatchGroup = 8;                       // lanes per problem
constexpr int kBatchPerWarp = 32 / kBatchGroup;      // problems a warp solves at once
constexpr int kBatchColsPerLane = 8;                 // kBatchMaxCols / kBatchGroup
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

// One warp per CTA (no block-level synchronisation anywhere); shared memory = 4 problems.
// kCpl = columns per lane = ceil(cols / 8): the scan and the price update are unrolled to it.
template <int kCpl> __global__ void __launch_bounds__(32) k_auction_batch(BatchParams P)
{
    extern __shared__ __align__(16) float smem[];
    const int lane = threadIdx.x;
    const int g = lane / kBatchGroup, gl = lane % kBatchGroup;
    const int n = P.rows * P.cols;
    const int n_pad = (n + 3) & ~3;
    // problem g starts 8 g floats later than dense packing would put it: with a row stride of 48
    // floats the four groups then read disjoint 8-bank windows whenever their rows have the same
    // parity (ncu: 70 % of the shared-load wavefronts were bank conflicts without it)
    const int p_stride = n_pad + kBatchGroup;
    const float* sA = smem + (size_t)g * p_stride;
    const unsigned all_rows = P.rows == 32 ? kFullMask : ((1u << P.rows) - 1u);

    for (long long base = (long long)blockIdx.x * kBatchPerWarp; base < P.batch;
         base += (long long)gridDim.x * kBatchPerWarp)
    {
        const int count = (int)min((long long)kBatchPerWarp, P.batch - base);
        const float* gsrc = P.a + base * n;
        if ((n & 3) == 0)
        {
            // the four problems are contiguous in HBM: streaming 128-bit copies
            const int vecs = n / 4;
            for (int q = 0; q < count; q++)
                for (int k = lane; k < vecs; k += 32)
                    reinterpret_cast<float4*>(smem + (size_t)q * p_stride)[k] = ld_stream4(gsrc + (size_t)q * n + 4 * k);
        }
        else
        {
            for (int q = 0; q < count; q++)
                for (int k = lane; k < n; k += 32)
                    smem[(size_t)q * p_stride + k] = __ldg(gsrc + (size_t)q * n + k);
        }
        __syncwarp();

        double p[kCpl]; // prices of columns gl + 8k (Auction.h:58)
        int own[kCpl];  // rows holding them
        int mc[4];                   // columns of rows gl + 8k
#pragma unroll
        for (int k = 0; k < kCpl; k++)
        {
            p[k] = 0.0;
            own[k] = -1;
        }
#pragma unroll
        for (int k = 0; k < 4; k++)
            mc[k] = -1;
        unsigned unassigned = g < count ? all_rows : 0u; // uniform inside a group
        unsigned todo = 0;
        int bids = 0;
        int fail = 0;

        while (__any_sync(kFullMask, unassigned != 0))
        {
            if (todo == 0)
                todo = unassigned; // next forward sweep (Auction.h:66-68, :220): rows in increasing order
            const bool act = todo != 0;
            const int i = act ? __ffs(todo) - 1 : 0;
            todo &= todo - 1;
            // findForward (Auction.h:124-152) over this lane's columns, in increasing column order
            // two independent accumulators (even / odd k) halve the dependent chain; merging them
            // keeps "lowest column wins ties" because top2_merge compares columns
            Top2<float> t = top2_empty<float>(), t1 = top2_empty<float>();
            const float* row = sA + i * P.cols;
#pragma unroll
            for (int k = 0; k < kCpl; k++)
            {
                const int c = gl + kBatchGroup * k;
                if (k < kCpl - 1 || c < P.cols)
                    top2_push((k & 1) ? t1 : t, row[c], p[k], c);
            }
            if (kCpl > 1)
                top2_merge(t, t1);
#pragma unroll
            for (int m = kBatchGroup / 2; m >= 1; m >>= 1)
            {
                const Top2<float> o = top2_shfl_xor(t, m);
                top2_merge(t, o);
            }
            const bool ok = act && t.j >= 0;
            const double bid = (double)t.a - t.w + P.eps; // Auction.h:250
            // lambda = 0 <= bid always holds here (Auction.h:257): assign, evict the owner
            const int jl = t.j & (kBatchGroup - 1), jk = t.j >> 3;
            int prev = -1;
            if (ok && gl == jl)
            {
#pragma unroll
                for (int k = 0; k < kCpl; k++)
                    if (k == jk)
                    {
                        prev = own[k];
                        p[k] = bid;
                        own[k] = i;
                    }
            }
            prev = __shfl_sync(kFullMask, prev, (lane & ~(kBatchGroup - 1)) | jl);
            if (ok)
            {
#pragma unroll
                for (int k = 0; k < 4; k++)
                {
                    if (gl + kBatchGroup * k == i)
                        mc[k] = t.j;
                    if (gl + kBatchGroup * k == prev)
                        mc[k] = -1; // Auction.h:270
                }
                unassigned &= ~(1u << i);
                if (prev >= 0)
                {
                    unassigned |= 1u << prev;
                    if (prev > i)
                        todo |= 1u << prev; // unlocked ahead of the sweep position: visited in this sweep
                }
                if (++bids > P.max_bids)
                {
                    fail = 3;
                    unassigned = 0;
                    todo = 0;
                }
            }
            else if (act)
            {
                fail = 2; // a row without an edge: the reference would never return
                unassigned = 0;
                todo = 0;
            }
        }
        if (fail && gl == 0)
            *P.status = fail;
        if (g < count)
        {
            int* out = P.row2col + (base + g) * P.rows;
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (gl + kBatchGroup * k < P.rows)
                    out[gl + kBatchGroup * k] = mc[k];
        }
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
