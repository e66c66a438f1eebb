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

// ---------------------------------------------------------------------------------------------
// Streaming form: no shared memory at all.  A row of a small problem is read from global memory when
// it bids (the eight lanes of a group read eight consecutive floats per k: 32-byte sectors), so the
// first sweep reads every row exactly once from HBM, in a known order — the next row of the sweep is
// prefetched into registers while the current bid is computed — and the ~30 % re-bids hit L1/L2.
// Without the 24 KB of shared memory per warp the SM holds 24-32 warps instead of 9, which is what
// this latency-bound kernel (long dependent fp64 / shuffle chains per bid) needs.  Same bids, same
// order, same result as k_auction_batch.
// ---------------------------------------------------------------------------------------------
// kG = lanes per problem (8: four problems per warp, 4: eight problems per warp and twice the columns
// per lane — fewer reduction stages and more independent work per lane, at more registers);
// kCpl = columns per lane = ceil(cols / kG).
template <int kCpl, int kG, int kWarpsPerSm>
__global__ void __launch_bounds__(32, kWarpsPerSm) k_auction_batch_stream(BatchParams P)
{
    constexpr int kPerWarp = 32 / kG;     // problems a warp solves at once
    constexpr int kRpl = kBatchMaxRows / kG; // rows per lane (assignment registers)
    const int lane = threadIdx.x;
    const int g = lane / kG, gl = lane % kG;
    const int n = P.rows * P.cols;
    const unsigned all_rows = P.rows == 32 ? kFullMask : ((1u << P.rows) - 1u);

    for (long long base = (long long)blockIdx.x * kPerWarp; base < P.batch;
         base += (long long)gridDim.x * kPerWarp)
    {
        const int count = (int)min((long long)kPerWarp, P.batch - base);
        const float* __restrict__ A = P.a + (base + (g < count ? g : 0)) * n; // this group's problem

        double p[kCpl]; // prices of columns gl + 8k (Auction.h:58)
        int own[kCpl];  // rows holding them
        int mc[kRpl];   // columns of rows gl + kG k
        float nx[kCpl]; // prefetched row nx_i
        int nx_i = -1;
#pragma unroll
        for (int k = 0; k < kCpl; k++)
        {
            p[k] = 0.0;
            own[k] = -1;
            nx[k] = 0.f;
        }
#pragma unroll
        for (int k = 0; k < kRpl; k++)
            mc[k] = -1;
        unsigned unassigned = g < count ? all_rows : 0u; // uniform inside a group
        unsigned todo = 0;
        int bids = 0;
        int fail = 0;
        if (unassigned)
        {
            // row 0 opens the first sweep
#pragma unroll
            for (int k = 0; k < kCpl; k++)
            {
                const int c = gl + kG * k;
                nx[k] = (k < kCpl - 1 || c < P.cols) ? __ldg(A + c) : 0.f;
            }
            nx_i = 0;
        }

        while (__any_sync(kFullMask, unassigned != 0))
        {
            if (todo == 0)
                todo = unassigned; // next forward sweep (Auction.h:66-68, :220): rows in increasing order
            const bool act = todo != 0;
            const int i = act ? __ffs(todo) - 1 : 0;
            todo &= todo - 1;
            float a[kCpl];
            if (act && nx_i == i)
            {
#pragma unroll
                for (int k = 0; k < kCpl; k++)
                    a[k] = nx[k];
            }
            else
            {
                const float* row = A + i * P.cols;
#pragma unroll
                for (int k = 0; k < kCpl; k++)
                {
                    const int c = gl + kG * k;
                    a[k] = (act && (k < kCpl - 1 || c < P.cols)) ? __ldg(row + c) : 0.f;
                }
            }
            // the next row of this sweep, if the sweep goes on (an eviction ahead of it is rare and
            // only costs a direct load)
            if (todo != 0)
            {
                const int i2 = __ffs(todo) - 1;
                const float* row2 = A + i2 * P.cols;
#pragma unroll
                for (int k = 0; k < kCpl; k++)
                {
                    const int c = gl + kG * k;
                    nx[k] = (k < kCpl - 1 || c < P.cols) ? __ldg(row2 + c) : 0.f;
                }
                nx_i = i2;
            }
            else
                nx_i = -1;
            // findForward (Auction.h:124-152) over this lane's columns, in increasing column order
            Top2<float> t = top2_empty<float>(), t1 = top2_empty<float>();
#pragma unroll
            for (int k = 0; k < kCpl; k++)
            {
                const int c = gl + kG * k;
                if (k < kCpl - 1 || c < P.cols)
                    top2_push((k & 1) ? t1 : t, a[k], p[k], c);
            }
            if (kCpl > 1)
                top2_merge(t, t1);
#pragma unroll
            for (int m = kG / 2; m >= 1; m >>= 1)
            {
                const Top2<float> o = top2_shfl_xor(t, m);
                top2_merge(t, o);
            }
            const bool ok = act && t.j >= 0;
            const double bid = (double)t.a - t.w + P.eps; // Auction.h:250
            const int jl = t.j % kG, jk = t.j / kG;
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
            prev = __shfl_sync(kFullMask, prev, (lane & ~(kG - 1)) | jl);
            if (ok)
            {
#pragma unroll
                for (int k = 0; k < kRpl; k++)
                {
                    if (gl + kG * k == i)
                        mc[k] = t.j;
                    if (gl + kG * k == prev)
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
            for (int k = 0; k < kRpl; k++)
                if (gl + kG * k < P.rows)
                    out[gl + kG * k] = mc[k];
        }
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
