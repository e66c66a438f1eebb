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

// Device-side control block of a batched launch (zeroed by the host before every launch).
struct BatchCtrl
{
    unsigned long long next; // dynamic hand-out: the first problem no warp has reserved yet
    int status;              // b200lsap_status of a failed problem, 0 when none failed
    int takeovers;           // problems that reached the bid cap and were finished by a whole warp
    int max_bids;            // largest bid count of a problem in this launch
    int pad;
    unsigned long long hits, misses; // takeover bids answered from a row cache / by a full scan
    unsigned long long dbg[4];       // -DB200LSAP_TIMING: cycles of the takeover's sections
};

struct BatchParams
{
    const float* a;   // [batch][rows*cols], row-major inside a problem
    int* row2col;     // [batch][rows]
    long long batch;
    int rows, cols;
    double eps;       // eps_multiplier / cols (Auction.h:55)
    int max_bids;     // watchdog per problem
    int* status;      // device flag, set to a b200lsap_status on failure (= &ctl->status)
    BatchCtrl* ctl;   // filtered kernel: hand-out counter and statistics
    int cap;          // filtered kernel: a problem with this many bids is finished by its whole warp
    int chunk;        // filtered kernel: problems per reservation of the hand-out counter
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

// ---------------------------------------------------------------------------------------------
// Filtered form (the default for cols % 4 == 0): the same Gauss-Seidel bids in the same order, but a
// bid no longer evaluates every column in fp64.  findForward (Auction.h:124-152) needs three exact
// things: the best column (lowest column among equals), its coefficient, and the second-best value.
// A single-precision pass finds the three largest a_ij - (float)p_j of the row with the column number
// carried in the six low mantissa bits of each value (top-3 by value only, two elements at a time:
// four FMNMX/FMNMX3 per element, no index selects); the two leaders are then evaluated exactly in fp64,
// and the result is accepted only if the third key, raised by a bound on everything the fp32 pass may
// have lost (key truncation 2^-17, the subtraction's rounding 2^-24, the price's rounding 2^-24 * max
// price), still lies strictly below the exact second-best value.  Otherwise (about one bid in a
// thousand on uniform data; always when a row has fewer than two edges) the group re-scans the row
// with the exact fp64 top-2 of the other kernels, so every bid is the reference's bid whatever the
// data — the filter only decides how much work it takes.  "a_ij == 0 is no edge" (Auction.h:134) costs
// two instructions per element, so the pass starts without it and a zero among the two leaders sends the
// bid to the exact scan and switches the warp to the zero-aware pass for the rest of its slice.
// Coefficients are assumed finite (the reference's contract is [0, 1], LSAP.cpp:25).
//
// Layout: 2 lanes per problem, 16 problems per warp; lane gl of a group owns the float4 chunks
// 2m + gl of a row (columns 8m + 4gl .. +3), so a pair reads one 32-byte sector per instruction.
// The fp32 roundings of the prices, the column owners, two row buffers and the fp64 prices live in shared
// memory, 880 bytes per 32x48 problem: a bid updates one entry of each with a store instead of a select
// chain over registers.  The fp64 price is kept per ROW (the price of the column that row holds: a priced
// column always has an owner, so 32 slots cover any number of columns), the assignment is read back from
// the owners when a problem finishes, and the arrays are unpadded — 56 320 bytes per 4-warp CTA, four CTAs
// = 16 warps per SM, which is what this latency-bound kernel needs most (12 warps with padded arrays and
// per-column fp64 prices: 452 M problems/s; 16 warps: 470).  The next row of the sweep arrives by cp.async
// while the current one bids.  A warp owns a contiguous slice of the batch and hands the next problems to the groups that
// have finished (the bid count of a 32x48 problem ranges from 32 to 200+; a warp that waited for its
// slowest group would idle 27 % of its lanes); it does so when two groups are waiting, so the divergent
// hand-over code runs half as often.
// ---------------------------------------------------------------------------------------------
// key of "no edge here": finite (the column bits must not turn it into a NaN), its six low bits clear
// (an untouched accumulator then names column 0, a valid index into the shared arrays)
#define B200LSAP_FILTER_NONE __uint_as_float(0xFF61B1C0u) /* -3.0e38f */

struct Top3f
{
    float m1, m2, m3;
};

// two new elements into the sorted triple
__device__ __forceinline__ void top3_push2(Top3f& t, float x, float y)
{
    const float hi = fmaxf(x, y), lo = fminf(x, y);
    const float a = fminf(t.m1, hi), b = fminf(t.m2, hi), c = fminf(t.m1, lo);
    t.m3 = fmaxf(fmaxf(t.m3, b), c);
    t.m2 = fmaxf(fmaxf(t.m2, a), lo);
    t.m1 = fmaxf(t.m1, hi);
}

// both sorted; result = the three largest of the six
__device__ __forceinline__ void top3_merge(Top3f& t, const Top3f& o)
{
    const float x = fminf(t.m1, o.m1), y = fmaxf(t.m2, o.m2);
    t.m3 = fmaxf(fmaxf(fminf(x, y), t.m3), o.m3);
    t.m2 = fmaxf(x, y);
    t.m1 = fmaxf(t.m1, o.m1);
}

template <bool kZeroAware> __device__ __forceinline__ float filter_key(float a, float p32, int col)
{
    const float d = (!kZeroAware || a != 0.f) ? a - p32 : B200LSAP_FILTER_NONE;
    return __uint_as_float((__float_as_uint(d) & ~63u) | (unsigned)col);
}

// The arrays are unpadded, so the problems of a quarter warp start 8 kM floats apart: for kM = 2 mod 4 two of
// the four share a 32-byte bank window, for kM = 0 mod 4 all four do.  The groups therefore walk their chunks in
// rotated order (group g starts at chunk rot(g)): at every step the four 128-bit accesses of a quarter warp
// then fall into different windows.  The order is free in the single-precision pass (keys are distinct) and
// in the copies; the exact scan keeps increasing columns.
template <int kM> __device__ __forceinline__ int filter_rot(int g)
{
    return kM % 4 == 0 ? (g & 3) : (kM % 4 == 2 ? ((g >> 1) & 1) : 0);
}
// first column of the chunk a lane visits at step m
template <int kM> __device__ __forceinline__ int filter_chunk(int m, int rot, int gl)
{
    constexpr int kRotMax = kM % 4 == 0 ? 3 : (kM % 4 == 2 ? 1 : 0);
    int mm = m + rot;
    if (m + kRotMax >= kM && mm >= kM) // only the last kRotMax steps can wrap
        mm -= kM;
    return 8 * mm + 4 * gl;
}

// the same relative to the lane's first chunk (column 8 rot + 4 gl): a constant for the steps that cannot wrap, so
// pointers formed once per row as base + first chunk take immediate offsets
template <int kM> __device__ __forceinline__ int filter_chunk_rel(int m, int rot)
{
    constexpr int kRotMax = kM % 4 == 0 ? 3 : (kM % 4 == 2 ? 1 : 0);
    if (m + kRotMax >= kM && m + rot >= kM)
        return 8 * (m - kM);
    return 8 * m;
}

template <bool kZeroAware, int kM>
__device__ __forceinline__ Top3f filter_pass(const float* __restrict__ row, const float* __restrict__ p32, int gl,
                                             int rot)
{
    // two independent triples (halves of every chunk) keep two dependent chains in flight
    Top3f t3 = {B200LSAP_FILTER_NONE, B200LSAP_FILTER_NONE, B200LSAP_FILTER_NONE}, u3 = t3;
#pragma unroll
    for (int m = 0; m < kM; m++)
    {
        const int c = filter_chunk<kM>(m, rot, gl);
        const float4 av = *reinterpret_cast<const float4*>(row + c);
        const float4 pp = *reinterpret_cast<const float4*>(p32 + c);
        top3_push2(t3, filter_key<kZeroAware>(av.x, pp.x, c), filter_key<kZeroAware>(av.y, pp.y, c + 1));
        top3_push2(u3, filter_key<kZeroAware>(av.z, pp.z, c + 2), filter_key<kZeroAware>(av.w, pp.w, c + 3));
    }
    top3_merge(t3, u3);
    return t3;
}

// stride (floats) of the per-problem fp32 arrays
__host__ __device__ inline int batch_filter_stride(int kM)
{
    return 8 * kM; // unpadded: a quarter warp's 128-bit accesses then take two wavefronts, 56 bytes per problem saved
}
// per warp: fp32 prices + two row buffers (3 x 16 x stride floats), fp64 prices by row, owners
__host__ __device__ inline size_t batch_filter_warp_bytes(int kM)
{
    return (size_t)16 * (3 * batch_filter_stride(kM) * 4 + kBatchMaxRows * 8 + 8 * kM);
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)),
                 "l"(gsrc)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kPending> __device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory");
}

// ---------------------------------------------------------------------------------------------
// Whole-warp takeover.  The bid count of a 32x48 problem has a long tail (mean 44; about 500 problems
// in a million need more than 96 bids, one or two need more than a thousand: price wars in which a few
// rows raise a few prices by eps per bid).  Two lanes play such a problem at ~0.7 us per bid while the
// rest of the launch has long finished, so a group that reaches P.cap bids hands its problem to its
// whole warp, which CONTINUES the group's Gauss-Seidel sweep (Auction.h:210-306) from its state
// (prices, owners, unassigned rows, position in the sweep): same bids in the same order, same result.
//
//  * The matrix goes into the warp's row buffers (exactly 32 x 8 kM floats); fp64 prices and owners live
//    per COLUMN in the group's own arrays.
//  * A full scan is the filter of the two-lane pass across the warp: lane l evaluates columns l and l + 32,
//    keys a - (float)p with the column in the six low bits, the largest keys one by one with redux.sync on
//    the order-preserving integer image of the keys, the two leaders evaluated exactly, the third key as
//    the certificate, and the exact fp64 top-2 of every column (shuffle merge) when the certificate fails.
//  * Row caches (the narrow rounds' idea, dense_cache.cuh, in miniature): a full scan of row r leaves the
//    columns of its four largest keys, their coefficients and an upper bound of a - p over every OTHER
//    column (the fifth key plus the filter's error margin) in the registers of lane r.  Prices never
//    decrease, so the bound holds for the rest of the solve.  The next bid of row r evaluates those four
//    columns exactly at the current prices — lane r alone, ~40 dependent instructions instead of ~100 —
//    and is the reference's bid whenever its second-best value lies strictly above the bound (nothing
//    outside the cache can then be best or second best); otherwise the row is scanned again.  A price war
//    is rows re-bidding for the same few columns: almost all of its bids hit.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned ord_u32(float x)
{
    const unsigned b = __float_as_uint(x);
    return b ^ ((unsigned)((int)b >> 31) | 0x80000000u); // larger float <=> larger unsigned
}
__device__ __forceinline__ float ord_f32(unsigned k)
{
    return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}

// what the single-precision keys may have lost against the exact a - p (two-lane pass: same bound)
__device__ __forceinline__ float filter_bound(float k, float pmax)
{
    return k + (fabsf(k) * 3.0517578125e-5f + pmax * 4.76837158203125e-7f + 1e-30f);
}

// Byte offsets into the kernel's dynamic shared memory: rows_off = the warp's row buffers (>= rows * cols floats,
// + 64 floats of the warp's own arrays behind them), own_off / p64_off / p32_off = the group's owner, fp64 price
// (by row) and fp32 price arrays; unassigned / todo / bids / pmax: the group's sweep state.  Writes the problem's
// row -> column (or -1 everywhere and *status on failure); returns the final bid count.
__device__ __noinline__ int batch_warp_finish(unsigned rows_off, unsigned own_off, unsigned p64_off, unsigned p32_off,
                                              const float* __restrict__ A, int* __restrict__ out, int rows, int cols,
                                              double eps, int max_bids, BatchCtrl* ctl, unsigned unassigned, unsigned todo,
                                              int bids, float pmax)
{
    extern __shared__ __align__(16) unsigned char batch_smem[];
    int* status = &ctl->status;
    float* sA = reinterpret_cast<float*>(batch_smem + rows_off);
    signed char* ownc = reinterpret_cast<signed char*>(batch_smem + own_off);   // owner of a column (stays as it is)
    double* pc_lo = reinterpret_cast<double*>(batch_smem + p64_off);            // price of columns 0..31
    double* pc_hi = reinterpret_cast<double*>(batch_smem + p32_off) - 32;       // price of columns 32..
    auto pc = [&](int c) -> double* { return (c < 32 ? pc_lo : pc_hi) + c; };
    const int lane = threadIdx.x & 31;
    const int n4 = rows * cols / 4;
    for (int k = lane; k < n4; k += 32)
        reinterpret_cast<float4*>(sA)[k] = __ldg(reinterpret_cast<const float4*>(A) + k);
    const int cA = lane, cB = lane + 32;
    const bool hasA = cA < cols, hasB = cB < cols;
    {
        // prices by row (the two-lane pass) -> prices by column (Auction.h:58), in place
        const int oA = hasA ? (int)ownc[cA] : -1, oB = hasB ? (int)ownc[cB] : -1;
        const double pA = oA >= 0 ? pc_lo[oA] : 0.0, pB = oB >= 0 ? pc_lo[oB] : 0.0;
        __syncwarp();
        if (hasA)
            pc_lo[cA] = pA;
        if (hasB)
            pc_hi[cB] = pB;
    }
    // this lane's row: cached columns (ascending, one per byte), their coefficients, bound of everything else
    unsigned cc = 0, pa01 = (p64_off >> 3) * 0x10001u, pa23 = pa01; // double index (a CTA's arrays may exceed 64 KB)
    double ca0 = 0.0, ca1 = 0.0, ca2 = 0.0, ca3 = 0.0;
    double cbound = CUDART_INF; // no cache yet: every bid misses
#ifdef B200LSAP_TIMING
    unsigned bt_acc[4] = {0, 0, 0, 0}, bt_t = clock();
#define BT_TICK(k, ...)                                                                                            \
    do                                                                                                             \
    {                                                                                                              \
        asm volatile("" ::__VA_ARGS__ : "memory");                                                                 \
        const unsigned now_ = clock();                                                                             \
        bt_acc[k] += now_ - bt_t;                                                                                  \
        bt_t = now_;                                                                                               \
    } while (0)
#else
#define BT_TICK(k, ...)
#endif
    int fail = 0, n_miss = 0;
    const int bids_in = bids;
    while (unassigned != 0)
    {
        if (todo == 0)
            todo = unassigned; // next forward sweep (Auction.h:66-68, :220)
        const int i = __ffs(todo) - 1;
        todo &= todo - 1;
        BT_TICK(3, "r"(i));
        __syncwarp(); // the previous bid's price and owner are visible
        // every lane evaluates its own row's cache (one instruction stream); lane i's answer is the bid
        int j, prev;
        double bid;
        {
            // (shared-memory indices of the cached columns' prices were formed when the cache was built.  Measured and
            // dropped: 32-bit shared addresses in four registers with ld.shared through inline asm, 0.20 -> 0.24 us per bid)
            const double* sm64 = reinterpret_cast<const double*>(batch_smem);
            const double d0 = ca0 - sm64[pa01 & 0xffffu], d1 = ca1 - sm64[pa01 >> 16], d2 = ca2 - sm64[pa23 & 0xffffu],
                         d3 = ca3 - sm64[pa23 >> 16];
            const int c0 = cc & 63, c1 = (cc >> 8) & 63, c2 = (cc >> 16) & 63, c3 = cc >> 24;
            const int o0 = ownc[c0], o1 = ownc[c1], o2 = ownc[c2], o3 = ownc[c3];
            // strict '>' in ascending column order keeps the lowest column among equals (Auction.h:136)
            const bool g01 = d1 > d0, g23 = d3 > d2;
            const double vA = g01 ? d1 : d0, wA = g01 ? d0 : d1, vB = g23 ? d3 : d2, wB = g23 ? d2 : d3;
            const bool gB = vB > vA;
            const double x2 = gB ? vA : vB, y2 = gB ? wB : wA;
            const double w = x2 > y2 ? x2 : y2; // second best, may equal the best (Auction.h:147-148)
            const double aj = gB ? (g23 ? ca3 : ca2) : (g01 ? ca1 : ca0);
            const int mj = gB ? (g23 ? c3 : c2) : (g01 ? c1 : c0);
            const int mo = gB ? (g23 ? o3 : o2) : (g01 ? o1 : o0);
            const bool hit = w > cbound;
            BT_TICK(0, "d"(w), "r"(mj), "r"(mo));
            j = __shfl_sync(kFullMask, hit ? mj : -1, i);
            prev = __shfl_sync(kFullMask, mo, i);
            bid = __shfl_sync(kFullMask, aj - w + eps, i); // Auction.h:250
            BT_TICK(1, "d"(bid), "r"(j), "r"(prev));
        }
        if (j < 0)
        {
            n_miss++;
            // full scan of row i, every lane two columns
            const float* row = sA + i * cols;
            const float a0 = hasA ? row[cA] : 0.f, a1 = hasB ? row[cB] : 0.f; // 0: no edge (Auction.h:134)
            const double pA = hasA ? pc_lo[cA] : 0.0, pB = hasB ? pc_hi[cB] : 0.0;
            const float fpA = __double2float_rn(pA), fpB = __double2float_rn(pB);
            // upper bound of the prices for the slack of the certificate: taken here, where every price is at hand, and
            // not tracked bid by bid (non-negative floats order like their bits)
            pmax = fmaxf(pmax, __uint_as_float(__reduce_max_sync(kFullMask, __float_as_uint(fmaxf(fmaxf(fpA, fpB), 0.f)))));
            const unsigned uA = ord_u32(filter_key<true>(a0, fpA, cA)), uB = ord_u32(filter_key<true>(a1, fpB, cB));
            // keys are distinct (the column is part of the key): each redux removes exactly one of them
            unsigned x = max(uA, uB), y = min(uA, uB);
            unsigned u[5];
#pragma unroll
            for (int t = 0; t < 5; t++)
            {
                u[t] = __reduce_max_sync(kFullMask, x);
                if (x == u[t])
                {
                    x = y;
                    y = 0u;
                }
            }
            const float k1 = ord_f32(u[0]), k2 = ord_f32(u[1]), k3 = ord_f32(u[2]), k4 = ord_f32(u[3]), k5 = ord_f32(u[4]);
            // the two leaders, exactly (Auction.h:131-150: strict '>' keeps the lowest column among equals)
            int c1 = __float_as_uint(k1) & 63, c2 = __float_as_uint(k2) & 63;
            const float e1 = row[c1], e2 = row[c2];
            const double d1 = (double)e1 - *pc(c1), d2 = (double)e2 - *pc(c2);
            const bool first = d1 > d2 || (d1 == d2 && c1 < c2);
            j = first ? c1 : c2;
            float aj = first ? e1 : e2;
            double wv = first ? d2 : d1;
            // nothing else can reach the second-best value
            const bool fast = k2 > -5.0e5f && wv > (double)filter_bound(k3, pmax) && e1 != 0.f && e2 != 0.f;
            if (!fast)
            {
                Top2<float> t = top2_empty<float>();
                top2_push(t, a0, pA, cA);
                top2_push(t, a1, pB, cB);
                top2_warp_reduce(t);
                j = t.j;
                aj = t.a;
                wv = t.w;
            }
            if (j < 0)
            {
                fail = 2; // a row without an edge: the reference would never return
                break;
            }
            bid = (double)aj - wv + eps; // Auction.h:250
            prev = ownc[j];
            // row i's cache: the columns of the four largest keys in ascending order; every other column's exact
            // a - p lies below the fifth key plus the filter's margin, now and for ever
            int c3 = __float_as_uint(k3) & 63, c4 = __float_as_uint(k4) & 63;
            int lo, hi;
#define B200LSAP_CSWAP(a, b) lo = min(a, b), hi = max(a, b), a = lo, b = hi
            B200LSAP_CSWAP(c1, c2);
            B200LSAP_CSWAP(c3, c4);
            B200LSAP_CSWAP(c1, c3);
            B200LSAP_CSWAP(c2, c4);
            B200LSAP_CSWAP(c2, c3);
#undef B200LSAP_CSWAP
            // (a key of a column beyond the row belongs to no edge: cached as column 0 with no coefficient)
            const bool in1 = c1 < cols, in2 = c2 < cols, in3 = c3 < cols, in4 = c4 < cols;
            c1 = in1 ? c1 : 0, c2 = in2 ? c2 : 0, c3 = in3 ? c3 : 0, c4 = in4 ? c4 : 0;
            const float f1 = in1 ? row[c1] : 0.f, f2 = in2 ? row[c2] : 0.f, f3 = in3 ? row[c3] : 0.f,
                        f4 = in4 ? row[c4] : 0.f;
            if (lane == i)
            {
                cc = (unsigned)c1 | ((unsigned)c2 << 8) | ((unsigned)c3 << 16) | ((unsigned)c4 << 24);
                const unsigned lo_idx = p64_off >> 3, hi_idx = (p32_off >> 3) - 32u;
                pa01 = ((c1 < 32 ? lo_idx : hi_idx) + c1) | (((c2 < 32 ? lo_idx : hi_idx) + c2) << 16);
                pa23 = ((c3 < 32 ? lo_idx : hi_idx) + c3) | (((c4 < 32 ? lo_idx : hi_idx) + c4) << 16);
                // a column without an edge can neither win nor be second: its value sinks below every bound
                ca0 = f1 != 0.f ? (double)f1 : -1e300;
                ca1 = f2 != 0.f ? (double)f2 : -1e300;
                ca2 = f3 != 0.f ? (double)f3 : -1e300;
                ca3 = f4 != 0.f ? (double)f4 : -1e300;
                cbound = (double)filter_bound(k5, pmax);
            }
        }
        if (lane == i)
        {
            *pc(j) = bid;
            ownc[j] = (signed char)i;
        }
        unassigned &= ~(1u << i);
        if (prev >= 0)
        {
            unassigned |= 1u << prev; // Auction.h:270
            if (prev > i)
                todo |= 1u << prev; // unlocked ahead of the sweep position: visited in this sweep
        }
        if (++bids > max_bids)
        {
            fail = 3;
            break;
        }
        BT_TICK(2, "r"(unassigned), "r"(todo));
    }
#ifdef B200LSAP_TIMING
    if (lane == 0)
        for (int k = 0; k < 4; k++)
            atomicAdd(&ctl->dbg[k], (unsigned long long)bt_acc[k]);
#endif
    __syncwarp();
    if (fail)
    {
        for (int r = lane; r < rows; r += 32)
            out[r] = -1;
        if (lane == 0)
            *status = fail;
    }
    else
    {
        const int oA = hasA ? (int)ownc[cA] : -1, oB = hasB ? (int)ownc[cB] : -1;
        if (oA >= 0)
            out[oA] = cA;
        if (oB >= 0)
            out[oB] = cB;
    }
    if (lane == 0)
    {
        atomicAdd(&ctl->misses, (unsigned long long)n_miss);
        atomicAdd(&ctl->hits, (unsigned long long)(bids - bids_in - n_miss));
    }
    __syncwarp();
    return bids;
}

// kWhole: cols % 8 == 0, every lane's last chunk lies inside the row (no predicate around its copies)
template <int kM, bool kWhole, int kWarps, int kCtasPerSm>
__global__ void __launch_bounds__(kWarps * 32, kCtasPerSm) k_auction_batch_filter(BatchParams P)
{
    extern __shared__ __align__(16) unsigned char batch_smem[];
    constexpr int kC8 = 8 * kM;
    const int s32 = batch_filter_stride(kM);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 1, gl = lane & 1;
    unsigned char* wb = batch_smem + warp * batch_filter_warp_bytes(kM);
    float* p32 = reinterpret_cast<float*>(wb) + g * s32;                       // (float)price[col]
    float* rowbuf = reinterpret_cast<float*>(wb) + 16 * s32 + g * s32;         // + cur * 16 * s32: the bidding row
    // fp64 price (Auction.h:58) of the column a row holds, indexed by the ROW: at most `rows` columns are priced
    // at any time (a priced column always has an owner), a column without an owner costs 0
    double* p64r = reinterpret_cast<double*>(wb + 3 * 16 * s32 * 4) + g * kBatchMaxRows;
    signed char* own = reinterpret_cast<signed char*>(wb + 3 * 16 * s32 * 4 + 16 * kBatchMaxRows * 8) + g * kC8;
    auto price = [&](int c) -> double {
        const int o = own[c];
        const double q = p64r[o & 31]; // a group without a problem reads whatever its shared memory holds
        return o >= 0 ? q : 0.0;
    };
    const int n = P.rows * P.cols;
    const unsigned all_rows = P.rows == 32 ? kFullMask : ((1u << P.rows) - 1u);
    const unsigned lanes_before = (1u << lane) - 1u;
    const int rot = filter_rot<kM>(g);
    // this lane's last chunk lies beyond the row when cols % 8 == 4: never loaded, reads as "no edge"
    const bool last_chunk = kWhole || 8 * (kM - 1) + 4 * gl < P.cols;
    if (!last_chunk)
    {
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(rowbuf + 8 * (kM - 1) + 4 * gl) = z;
        *reinterpret_cast<float4*>(rowbuf + 16 * s32 + 8 * (kM - 1) + 4 * gl) = z;
    }

    // Dynamic hand-out: a warp reserves P.chunk problems at a time from one counter, so nobody owns a slice whose
    // slow problems it must play alone; the reservation after the current one is requested a reservation ahead
    // (lane 0 holds it: the atomic's latency is spent while the current problems bid).
    long long next, end, nbase = 0;
    {
        unsigned long long b0 = 0;
        if (lane == 0)
        {
            b0 = atomicAdd(&P.ctl->next, (unsigned long long)P.chunk);
            nbase = b0 < (unsigned long long)P.batch ? (long long)atomicAdd(&P.ctl->next, (unsigned long long)P.chunk)
                                                     : (long long)b0;
        }
        next = (long long)__shfl_sync(kFullMask, b0, 0);
        end = min(next + P.chunk, P.batch);
    }
    int max_seen = 0;               // largest bid count among this lane's problems

    long long pb = -1;              // the group's problem
    const float* __restrict__ A = P.a;
    unsigned unassigned = 0, todo = 0;
    int bids = 0, fail = 0;
    bool done = false;
    bool zero_aware = false;        // warp-uniform, sticky
    float pmax = 0.f;               // upper bound of the problem's prices
    int cur = 0;                    // row buffer of the current bid; the other one receives the prefetch
    int pf_i = -1;                  // row the other buffer holds (or will, once its cp.async lands)

    while (true)
    {
        const bool need = unassigned == 0 && !done;
        const unsigned nm = __ballot_sync(kFullMask, need);
        const unsigned am = __ballot_sync(kFullMask, unassigned != 0);
        if (__popc(nm) >= 4 || (nm != 0 && am == 0))
        {
            const int k = __popc(nm) >> 1;                 // groups that want a problem
            const int r = __popc(nm & lanes_before) >> 1;  // this group's rank among them
            const long long avail = max(end - next, 0ll);
            const bool sw = avail < k;                     // the current reservation does not cover them: open the next
            long long nb = 0, nend = 0;
            if (sw)
            {
                nb = __shfl_sync(kFullMask, nbase, 0);
                nend = min(nb + P.chunk, P.batch);
            }
            if (need)
            {
                if (pb >= 0)
                {
                    // row -> column from the owners (a failed problem: -1 everywhere)
                    int* out = P.row2col + pb * P.rows;
                    if (fail)
                    {
                        for (int rr = gl; rr < P.rows; rr += 2)
                            out[rr] = -1;
                    }
                    else
                    {
#pragma unroll
                        for (int m = 0; m < kM; m++)
                        {
                            const int q = *reinterpret_cast<const int*>(own + 8 * m + 4 * gl);
#pragma unroll
                            for (int e = 0; e < 4; e++)
                            {
                                const int o = (int)(signed char)(q >> (8 * e));
                                if (o >= 0)
                                    out[o] = 8 * m + 4 * gl + e;
                            }
                        }
                    }
                    if (fail && gl == 0)
                        *P.status = fail;
                }
                max_seen = max(max_seen, bids);
                bids = 0;
                long long np = -1;
                if (r < avail)
                    np = next + r;
                else if (sw && nb + (r - avail) < nend)
                    np = nb + (r - avail);
                if (np >= 0)
                {
                    pb = np;
                    A = P.a + pb * n;
                    float* nbuf = rowbuf + (cur ^ 1) * 16 * s32;
#pragma unroll
                    for (int m = 0; m < kM; m++)
                    {
                        const int c = 8 * m + 4 * gl;
                        // a column beyond the row is priced out of the single-precision pass
                        const float p0 = (m < kM - 1 || last_chunk) ? 0.f : 3.0e38f;
                        *reinterpret_cast<float4*>(p32 + c) = make_float4(p0, p0, p0, p0);
                        *reinterpret_cast<int*>(own + c) = -1;
                        if (m < kM - 1 || last_chunk)
                            cp_async16(nbuf + c, A + c);
                    }
                    pf_i = 0;
                    unassigned = all_rows;
                    todo = 0;
                    bids = 0;
                    fail = 0;
                    pmax = 0.f;
                }
                else
                {
                    pb = -1;            // nothing in reach this time; nothing left at all once the counter has passed the batch
                    done = sw && nb >= P.batch;
                }
            }
            if (sw)
            {
                next = nb + max(0ll, min((long long)k - avail, nend - nb));
                end = nend;
                if (lane == 0)
                    nbase = nb < P.batch ? (long long)atomicAdd(&P.ctl->next, (unsigned long long)P.chunk) : nb;
            }
            else
                next += k;
            __syncwarp();
            if (__all_sync(kFullMask, done))
                break;
        }

        // a problem that has reached the bid cap is finished by the whole warp (batch_warp_finish).  (Also handing over
        // a warp's last one or two problems once the batch has run dry was measured: 2 % slower, the takeovers' cold
        // caches cost more than the idle lanes.)
        const unsigned cm = __ballot_sync(kFullMask, unassigned != 0 && bids >= P.cap);
        if (cm != 0)
        {
            const int src = __ffs(cm) - 1;
            const long long t_pb = __shfl_sync(kFullMask, pb, src);
            const unsigned t_un = __shfl_sync(kFullMask, unassigned, src), t_todo = __shfl_sync(kFullMask, todo, src);
            const int t_bids = __shfl_sync(kFullMask, bids, src);
            const float t_pmax = __shfl_sync(kFullMask, pmax, src);
            cp_async_wait<0>(); // the row buffers are about to hold the whole matrix
            __syncwarp();
            const int tg = src >> 1;
            const unsigned wb_off = warp * (unsigned)batch_filter_warp_bytes(kM);
            const int total = batch_warp_finish(
                wb_off + 16 * s32 * 4, wb_off + 3 * 16 * s32 * 4 + 16 * kBatchMaxRows * 8 + tg * kC8,
                wb_off + 3 * 16 * s32 * 4 + tg * kBatchMaxRows * 8, wb_off + tg * s32 * 4, P.a + t_pb * n,
                P.row2col + t_pb * P.rows, P.rows, P.cols, P.eps, P.max_bids, P.ctl, t_un, t_todo, t_bids, t_pmax);
            if (lane == 0)
                atomicAdd(&P.ctl->takeovers, 1);
            max_seen = max(max_seen, total);
            if (g == tg)
            {
                pb = -1; // already written
                unassigned = 0;
                todo = 0;
                bids = 0;
            }
            pf_i = -1; // every group's prefetched row is gone
            if (!last_chunk)
            {
                const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                *reinterpret_cast<float4*>(rowbuf + 8 * (kM - 1) + 4 * gl) = z;
                *reinterpret_cast<float4*>(rowbuf + 16 * s32 + 8 * (kM - 1) + 4 * gl) = z;
            }
            __syncwarp();
            continue;
        }

        if (todo == 0)
            todo = unassigned; // next forward sweep (Auction.h:66-68, :220): rows in increasing order
        const bool act = todo != 0;
        const int i = act ? __ffs(todo) - 1 : 0;
        todo &= todo - 1;
        // the row of this bid: normally the one prefetched during the previous bid
        if (act)
        {
            cur ^= 1;
            if (pf_i != i)
            {
                // this buffer may still be receiving the row that was prefetched instead (issued a bid ago):
                // it must have landed before the right row is copied over it
                cp_async_wait<0>();
                const float* row = A + i * P.cols;
                float* cb = rowbuf + cur * 16 * s32;
#pragma unroll
                for (int m = 0; m < kM; m++)
                {
                    const int c = filter_chunk<kM>(m, rot, gl);
                    if (c < 8 * (kM - 1) || last_chunk)
                        cp_async16(cb + c, row + c);
                }
            }
        }
        cp_async_commit();
        // the next row of this sweep, if the sweep goes on (an eviction ahead of it is rare and only
        // costs the direct load above)
        // (after the last row of a sweep: the lowest row this sweep has unlocked so far, which opens the next
        // sweep unless this very bid unlocks a lower one — 1.8 fewer waits per 32x48 problem)
        const unsigned ahead = todo != 0 ? todo : (unassigned & ~(1u << i));
        if (ahead != 0)
        {
            pf_i = __ffs(ahead) - 1;
            const float* row2 = A + pf_i * P.cols + (8 * rot + 4 * gl); // this lane's first chunk
            float* nb = rowbuf + (cur ^ 1) * 16 * s32 + (8 * rot + 4 * gl);
#pragma unroll
            for (int m = 0; m < kM; m++)
            {
                const int rel = filter_chunk_rel<kM>(m, rot);
                if (kWhole || filter_chunk<kM>(m, rot, gl) < 8 * (kM - 1) || last_chunk)
                    cp_async16(nb + rel, row2 + rel);
            }
        }
        else
            pf_i = -1;
        cp_async_commit();
        cp_async_wait<1>();
        const float* row = rowbuf + cur * 16 * s32;

        // single-precision pass over this lane's chunks (its own copies: no synchronisation needed yet)
        Top3f t3 = zero_aware ? filter_pass<true, kM>(row, p32, gl, rot) : filter_pass<false, kM>(row, p32, gl, rot);
        {
            Top3f o;
            o.m1 = __shfl_xor_sync(kFullMask, t3.m1, 1);
            o.m2 = __shfl_xor_sync(kFullMask, t3.m2, 1);
            o.m3 = __shfl_xor_sync(kFullMask, t3.m3, 1);
            top3_merge(t3, o);
        }
        __syncwarp(); // the partner's half of the row has landed too

        // the two leaders, exactly (Auction.h:131-150: strict '>' keeps the lowest column among equals)
        const int c1 = __float_as_uint(t3.m1) & 63, c2 = __float_as_uint(t3.m2) & 63;
        const float a1 = row[c1], a2 = row[c2];
        const int o1 = own[c1], o2 = own[c2];
        const double q1 = p64r[o1 & 31], q2 = p64r[o2 & 31];
        const double d1 = (double)a1 - (o1 >= 0 ? q1 : 0.0), d2 = (double)a2 - (o2 >= 0 ? q2 : 0.0);
        const bool first = d1 > d2 || (d1 == d2 && c1 < c2);
        int j = first ? c1 : c2;
        float aj = first ? a1 : a2;
        double wv = first ? d2 : d1;
        // nothing else can reach the second-best value
        const float k3 = t3.m3;
        const float bound = k3 + (fabsf(k3) * 3.0517578125e-5f + pmax * 4.76837158203125e-7f + 1e-30f);
        const bool nonzero = a1 != 0.f && a2 != 0.f;
        const bool fast = t3.m2 > -5.0e5f && wv > (double)bound && nonzero;
        if (__any_sync(kFullMask, act && !fast))
        {
            if (__any_sync(kFullMask, act && !nonzero))
                zero_aware = true;
            // exact scan of the whole row (the other kernels' top-2), for the groups that need it
            Top2<float> t = top2_empty<float>();
#pragma unroll
            for (int m = 0; m < kM; m++)
            {
                const int c = 8 * m + 4 * gl;
                if (m < kM - 1 || last_chunk)
                {
                    const float4 av = *reinterpret_cast<const float4*>(row + c);
                    top2_push(t, av.x, price(c), c);
                    top2_push(t, av.y, price(c + 1), c + 1);
                    top2_push(t, av.z, price(c + 2), c + 2);
                    top2_push(t, av.w, price(c + 3), c + 3);
                }
            }
            const Top2<float> o = top2_shfl_xor(t, 1);
            top2_merge(t, o);
            if (!fast)
            {
                j = t.j;
                aj = t.a;
                wv = t.w;
            }
        }
        const bool ok = act && j >= 0;
        const double bid = (double)aj - wv + P.eps; // Auction.h:250
        // lambda = 0 <= bid always holds here (Auction.h:257): assign, evict the owner
        const int prev = ok ? (int)own[j] : -1;
        __syncwarp();
        if (ok)
        {
            const float b32 = __double2float_rn(bid);
            if (gl == 0)
            {
                own[j] = (signed char)i;
                p64r[i] = bid;
                p32[j] = b32;
            }
            pmax = fmaxf(pmax, b32);
            unassigned &= ~(1u << i);
            if (prev >= 0)
            {
                unassigned |= 1u << prev; // Auction.h:270
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
        __syncwarp();
    }
    cp_async_wait<0>();
    max_seen = max(max_seen, bids);
    max_seen = __reduce_max_sync(kFullMask, max_seen);
    if (lane == 0 && max_seen > 0)
        atomicMax(&P.ctl->max_bids, max_seen);
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
