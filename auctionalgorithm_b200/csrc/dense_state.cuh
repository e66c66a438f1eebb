// dense_state.cuh — device-side state of one dense auction and the layout kernels.
//
// HBM layout: the benefit matrix is held ROW-major, A[i*ld + j], ld a multiple of
// 32 elements, columns [cols, ld) zero-filled.  A zero is the reference's "no edge" value
// (Auction.h:134), so the padding is skipped by the same test that skips real zeros and the
// scan loops need no column tail.  Eigen's default column-major input is
// transposed once on upload.
#pragma once

#include "common.cuh"
#include "../../include/b200lsap_gen.h"

namespace b200lsap
{

constexpr int kMaxPhases = 48;
constexpr int kCand = 32;       // candidate (column, a) pairs kept per row: slot = (column / 4) % 32
constexpr int kNarrowMax = 256; // power of two >= the largest narrow round (SMs - 1 bidders)
constexpr int kCacheK = 32;     // row cache: the best (column, a) pairs of a row at its last CTA scan, one per lane
constexpr int kSurvCap = 256;   // most elements a CTA scan may pass on to the exact evaluation
constexpr int kProbeRows = 128; // rows sampled to size the first eps of the scaling schedule

// Row cache entry (dense_cache.cuh): a column of the row and its coefficient.  a == 0 ("no edge",
// Auction.h:134) marks an unused slot; every slot always names a real column.
template <typename T> struct CacheEnt;
template <> struct __align__(8) CacheEnt<float>
{
    int col;
    float a;
};
template <> struct __align__(16) CacheEnt<double>
{
    double a;
    int col;
    int pad;
};
template <typename T> __device__ __forceinline__ CacheEnt<T> make_ent(int col, T a)
{
    CacheEnt<T> e{};
    e.col = col;
    e.a = a;
    return e;
}

// Row cache header (dense_cache.cuh): `count` valid entries in cache[row][0..count); every element of
// the row that is NOT cached had a - p <= bound when the cache was built.  Prices never decrease,
// so that stays true, and a bid whose second best cached value exceeds `bound` is exact.
struct __align__(16) RowCacheHdr
{
    double bound;
    int count;
    int pad;
};

// Controller block, one per solve, in device memory (mirrored to pinned host memory at the end).
struct DenseCtrl
{
    int count[2];      // bidders in list[0] / list[1]
    int status;        // b200lsap_status raised on the device
    int phase;         // current eps phase
    int n_phases;
    int pad0;
    long long rounds;
    long long bids;
    long long max_rounds;
    long long narrow_rounds; // rounds run by the controller / scanner hand-shake (included in rounds)
    long long gs_bids;      // bids placed by the Gauss-Seidel tail (included in bids)
    long long gs_episodes;  // times the tail ran
    long long dbg[16];      // cycle counters of the narrow rounds / tail (B200LSAP_TIMING builds)
    long long cache_hits;   // narrow-round bids answered from the row cache alone
    long long cache_misses; // narrow-round bids that needed a CTA scan of the row
    long long cache_fallbacks; // CTA scans that fell back to the unfiltered path (ties, overflow, < 2 edges)
    int eps_auto;           // 1: k_schedule derives eps[0] from the probe (square problems, fp32)
    int cur;                // which of list[0] / list[1] holds the bidders (kept across the launches of a solve)
    int done;               // the last phase has finished
    int pad1;
    double eps_factor;
    double eps_final;
    double eps_probe_scale; // eps[0] = this fraction of the mean probe spread (default 1/8)
    // safeguard of the data-sized start (narrow_kernel.cuh, escalate_schedule): the first phase of the
    // schedule may spend phase_budget work units (rounds + 4 per cache miss); beyond that its eps was too
    // small for this input and the schedule restarts from eps * factor^2 (at most eps_cap), prices kept
    long long phase_work;   // work units of the current phase so far
    long long phase_budget; // 0: no safeguard (explicit eps_start, fixed eps)
    long long escalations;
    int esc_phase;          // index of the first phase of the current schedule
    int pad2;
    double eps_cap;
    double probe[kProbeRows]; // best minus cache bound of the sampled rows at the initial prices, < 0: no sample
    double pmax;            // upper bound of every price so far (monotone; bits ordered as u64)
    double amax;            // max |a(i,j)| of the matrix being solved (k_absmax)
    double eps[kMaxPhases]; // eps[n_phases-1] = eps_multiplier / cols
    unsigned long long shard_x; // fused sharded solve: exchanges so far (handed from CTA 0 back to the grid)
    // rows < cols with eps scaling or a warm start: price post-condition at the end of every phase (asym_repair)
    unsigned long long lam_hi;  // bits of the highest price of an unassigned object
    unsigned long long lam_lo;  // bits of the lowest price of an assigned object
    long long repairs;          // floor raises so far
    int solo_seq;               // episodes that one CTA has played alone so far (the others doze on this word)
    int cache_fresh;            // row caches: 1 = built in this solve and the last narrow phase hardly missed (they hold across
                                // phases: prices only rise, so every bound stays valid); 0 = (re)build before the next one
};

// Slack of the single-precision filter of scan_segment_filtered: with |a| <= amax, 0 <= p <= pbound,
// |fl32(a - fl32(p)) - fl64(a - p)| < 2^-23 (amax + pbound); 2^-21 (amax + pbound + 1) also covers
// the roundings of the threshold itself.  Rounded up.
__device__ __forceinline__ float filter_margin(double amax, double pbound)
{
    return __double2float_ru((amax + pbound + 1.0) * (1.0 / 2097152.0));
}

// Narrow rounds: one mailbox per scanner CTA, each on its own 128-byte line so that a spinning
// scanner loads a line nobody else touches (148 SMs polling ONE address slowed every L2 access of
// the controller several-fold).  The controller fills the request and then sets `flag`; the scanner
// answers in place and bumps NarrowCtrl::arrived.
struct __align__(128) NarrowSlot
{
    int flag;  // > 0: sequence number of the round to serve; < 0: leave narrow mode
    int row;   // bidder row
    int c0, c1; // columns to scan
    double pmax; // price bound for the filter margin
    // answer
    double v, w, a; // best value, second best value, coefficient of the best (widened exactly)
    int j;          // its column, -1: nothing found in this part
    int pad[15];
};

struct NarrowCtrl
{
    unsigned arrived; // monotone count of answers published by scanners
    int pad[31];
    NarrowSlot slot[kNarrowMax];
};

template <typename T> struct DenseState
{
    const T* A;
    long long ld;
    int rows;
    int cols;
    int cols4; // cols rounded up to 4 (<= ld): scan bound, padding is zero
    double* prices;
    float* pricesf;  // (float)prices, read by the single-precision filter of the fp32 scan
    double* profits;
    BidSlot* slots;  // [cols]
    int* owner;      // [cols] row holding the column, -1 free
    int* row2col;    // [rows] column held by the row, -1 unassigned
    int* bid_col;    // [rows] column the row bid for this round
    double* bid_val; // [rows] its bid
    int* list[2];    // compacted unassigned rows, double buffered
    Top2<T>* partial; // [capacity >= max(total warps, 4 rows)] split-row partial results
    int* arrive;     // [rows] split-row arrival counters (zero between rounds)
    DenseCtrl* ctrl;
    int min_seg;     // smallest row segment a warp takes when rows are split (multiple of 128)
    int sweep_split; // rounds with more bidders than warps: 0 balance the last wave, 1 whole rows, >= 2 parts per row
    int2* cand;      // [rows][kCand] (column, bits of a): edges that were among the row's best, -1: empty
    NarrowCtrl* narrow;
    int max_parts;   // narrow rounds: at most this many CTAs share one bidder's row (power of two)
    int narrow_rows; // a round with at most this many bidders (and at most SMs - 1) runs as a narrow round
    int tail_rows;   // a round with at most this many bidders is finished by the Gauss-Seidel tail (0: never)
    int dry;         // 1: measurement sweep, bids are computed and stored but not offered
    CacheEnt<T>* cache;    // [rows][kCacheK] (column, a), best first
    RowCacheHdr* cache_hdr; // [rows]
    int use_cache;         // 1: narrow rounds answer from the row cache (persistent driver)
    int rebuild_caches_every_phase; // 1: the round-1 behaviour (developer knob B200LSAP_CACHE_EVERY_PHASE)
    int asym_repair;       // R > 0 (the last R phases repair): rows < cols and prices may break "no unassigned object priced above an assigned one"
                           // (eps scaling, warm start): checked and repaired at the end of every eps phase
    int phase_cut;         // an eps phase other than the last ends as soon as at most this many rows are unassigned
                           // (the next phase drops every assignment anyway; only the last phase decides the result)
    int narrow_after;      // row caches: a phase plays this many of its small rounds as wide rounds before it pays for
                           // the cache-building sweep (a pass over the whole matrix: 4 ms at 65536 x 98304, where a
                           // random asymmetric problem has ~50 small rounds of ~20 us left); 0: hand over at once
};

// ---------------------------------------------------------------------------------------------
// Column-major [rows x cols] (ld_in) -> row-major (ld_out), zero padding columns [cols, ld_out).
// 32x32 tiles through shared memory, both sides coalesced.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_transpose_pad(const T* __restrict__ in, long long ld_in, T* __restrict__ out,
                                                       long long ld_out, int rows, int cols)
{
    __shared__ T tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5; // 32 x 8
    const long long tiles_r = (rows + 31) / 32;
    const long long tiles_c = (ld_out + 31) / 32;
    for (long long tix = blockIdx.x; tix < tiles_r * tiles_c; tix += gridDim.x)
    {
        const int r0 = (int)(tix % tiles_r) * 32, c0 = (int)(tix / tiles_r) * 32;
#pragma unroll
        for (int k = 0; k < 32; k += 8)
        {
            const int c = c0 + ty + k, r = r0 + tx;
            tile[ty + k][tx] = (c < cols && r < rows) ? in[(long long)c * ld_in + r] : T(0);
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 32; k += 8)
        {
            const int r = r0 + ty + k, c = c0 + tx;
            if (r < rows && c < ld_out)
                out[(long long)r * ld_out + c] = tile[tx][ty + k];
        }
        __syncthreads();
    }
}

// Row-major with an arbitrary ld_in -> padded row-major.
template <typename T>
__global__ void __launch_bounds__(256) k_copy_pad(const T* __restrict__ in, long long ld_in, T* __restrict__ out,
                                                  long long ld_out, int rows, int cols)
{
    const long long total = (long long)rows * ld_out;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x)
    {
        const long long r = e / ld_out;
        const int c = (int)(e - r * ld_out);
        out[e] = c < cols ? in[r * ld_in + c] : T(0);
    }
}

// max |a(i,j)| over the padded matrix -> ctrl->amax (non-negative doubles order like their bits).
// NaNs are ignored, as they are by every comparison of the reference's scan.
template <typename T>
__global__ void __launch_bounds__(256) k_absmax(const T* __restrict__ A, long long ld, int rows, int cols4,
                                                DenseCtrl* ctrl)
{
    using V4 = typename Vec4<T>::type;
    const long long vec_per_row = cols4 / 4;
    const long long total = (long long)rows * vec_per_row;
    T m = T(0);
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long e0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; e0 < total; e0 += 4 * step)
    {
        // four independent 128-bit loads in flight per thread (one at a time left the kernel at 1.5 TB/s)
        V4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++)
        {
            const long long e = e0 + u * step;
            v[u].x = v[u].y = v[u].z = v[u].w = T(0);
            if (e < total)
            {
                const long long r = e / vec_per_row, c = (e - r * vec_per_row) * 4;
                v[u] = ld_stream4(A + r * ld + c);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++)
            m = fmax(fmax(m, fabs(v[u].x)), fmax(fabs(v[u].y), fmax(fabs(v[u].z), fabs(v[u].w))));
    }
    double d = (double)m;
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1)
        d = fmax(d, __shfl_xor_sync(kFullMask, d, s));
    if ((threadIdx.x & 31) == 0 && d > 0.0)
        atomicMax(reinterpret_cast<unsigned long long*>(&ctrl->amax), (unsigned long long)__double_as_longlong(d));
}

// Synthetic inputs, bit-identical to oracle_generate_f32 (include/b200lsap_gen.h).
// Element (b, i, j) of a batch; col0/cols_global select a column shard of one big problem.
__global__ void __launch_bounds__(256) k_generate_f32(float* __restrict__ out, long long batch, long long rows,
                                                      long long cols, long long ld, int layout,
                                                      unsigned long long seed, unsigned long long problem0,
                                                      long long col0, long long cols_global)
{
    const long long minor = layout ? cols : rows;
    const long long major = layout ? rows : cols;
    const long long per_problem = ld * major;
    const long long total = batch * per_problem;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x)
    {
        const long long b = e / per_problem;
        const long long r = e - b * per_problem;
        const long long mj = r / ld, mn = r - mj * ld;
        float v = 0.f;
        if (mn < minor)
        {
            const long long i = layout ? mj : mn, j = layout ? mn : mj;
            const uint64_t key = b200lsap_gen_key(seed, problem0 + (unsigned long long)b);
            v = b200lsap_gen_value(key, (uint64_t)(i * cols_global + col0 + j));
        }
        out[e] = v;
    }
}

} // namespace b200lsap
