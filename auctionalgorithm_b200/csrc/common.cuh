// common.cuh — shared device helpers for the sm_100a auction kernels.
//
// Reference semantics mirrored here (paths relative to /root/reference/AuctionAlgorithm/):
//   Top2 / top2_push   Auction.h:131-150     strict '>' best, second best may tie the best,
//                                            a == 0 is "no edge", init v = w = -1e6 (internal.h:17)
//   top2_merge         AuctionMT.h:377-398   exact merge of per-interval results, lowest column
//                                            wins ties (thread 0 scans intervals in order)
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace b200lsap
{

constexpr double kNoSecondBest = 1e6; // __AUCTION_INF
constexpr unsigned kFullMask = 0xffffffffu;

// ---------------------------------------------------------------------------------------------
// Per-column bid slot: (bid as ordered u64, 0xFFFFFFFF - row).  The lexicographic maximum is
// "highest bid, then lowest row", resolved with one 128-bit compare-and-swap per attempt, so
// the winner of a round does not depend on the order in which warps arrive.  Bids are
// strictly positive doubles (bid >= price + eps > 0), whose bit patterns order like u64.
// A slot is never reset: every later bid on the column exceeds its current price, which is
// the previous winning bid, so stale contents always lose.
// ---------------------------------------------------------------------------------------------
struct __align__(16) BidSlot
{
    unsigned long long bid;
    unsigned long long key;
};

__device__ __forceinline__ BidSlot cas128(BidSlot* addr, BidSlot expect, BidSlot desired)
{
    BidSlot old;
    asm volatile("{\n\t"
                 ".reg .b128 c, v, o;\n\t"
                 "mov.b128 c, {%2, %3};\n\t"
                 "mov.b128 v, {%4, %5};\n\t"
                 "atom.global.cas.b128 o, [%6], c, v;\n\t"
                 "mov.b128 {%0, %1}, o;\n\t"
                 "}"
                 : "=l"(old.bid), "=l"(old.key)
                 : "l"(expect.bid), "l"(expect.key), "l"(desired.bid), "l"(desired.key), "l"(addr)
                 : "memory");
    return old;
}

__device__ __forceinline__ unsigned long long row_key(int row) { return 0xFFFFFFFFull - (unsigned)row; }

__device__ __forceinline__ void slot_offer(BidSlot* slot, double bid, int row)
{
    BidSlot mine{(unsigned long long)__double_as_longlong(bid), row_key(row)};
    BidSlot seen;
    seen.bid = *reinterpret_cast<volatile unsigned long long*>(&slot->bid);
    seen.key = *reinterpret_cast<volatile unsigned long long*>(&slot->key);
    while (mine.bid > seen.bid || (mine.bid == seen.bid && mine.key > seen.key))
    {
        const BidSlot got = cas128(slot, seen, mine);
        if (got.bid == seen.bid && got.key == seen.key)
            break;
        seen = got;
    }
}

// ---------------------------------------------------------------------------------------------
// Running best / second best of a_ij - p_j for one row.
// ---------------------------------------------------------------------------------------------
template <typename T> struct Top2
{
    double v; // best value
    double w; // second best value (may equal v)
    int j;    // column of the best, -1: nothing found
    T a;      // a(i, j)
};

// kDeep = false is the reference's initial state (Auction.h:237).  kDeep = true starts from
// -infinity instead: used only to re-scan a row whose every candidate lies below -1e6, a state
// in which the reference's findForward reports "nothing found" and solve() never returns.
template <typename T, bool kDeep = false> __device__ __forceinline__ Top2<T> top2_empty()
{
    Top2<T> t;
    t.v = kDeep ? -CUDART_INF : -kNoSecondBest;
    t.w = kDeep ? -CUDART_INF : -kNoSecondBest;
    t.j = -1;
    t.a = T(0);
    return t;
}

// One element, branch-free.  Elements of a lane arrive in increasing column order, so the
// strict '>' keeps the lowest column among equals, as Auction.h:136 does.
template <typename T> __device__ __forceinline__ void top2_push(Top2<T>& t, T a, double p, int col)
{
    const double d = (double)a - p;
    const bool ok = a != T(0);
    const bool gv = ok && d > t.v;
    const bool gw = ok && d > t.w;
    t.w = gv ? t.v : (gw ? d : t.w);
    t.v = gv ? d : t.v;
    t.j = gv ? col : t.j;
    t.a = gv ? a : t.a;
}

// Exact float -> double widening on the integer pipe.  cvt.f64.f32 runs on the quarter-rate
// conversion unit, which saturates long before anything else when every element of a small
// problem is widened for every bid (ncu: XU pipe > 100 % busy in the batched kernel).  A normal
// finite float widens by re-biasing the exponent (+896) and shifting the mantissa by 29 bits;
// zeros, denormals, infinities and NaNs take the conversion instruction out of line.
__device__ __noinline__ double widen_slow(float a) { return (double)a; }

__device__ __forceinline__ double widen_exact(float a)
{
    const unsigned u = __float_as_uint(a);
    if ((u << 1) - 0x01000000u >= 0xFE000000u)
        return widen_slow(a);
    const unsigned hi = (u & 0x80000000u) | (((u & 0x7FFFFFFFu) >> 3) + 0x38000000u);
    return __hiloint2double((int)hi, (int)(u << 29));
}

// top2_push for fp32 coefficients with the widening done on the integer pipe (bit-identical).
__device__ __forceinline__ void top2_push_w(Top2<float>& t, float a, double p, int col)
{
    if (a == 0.f)
        return;
    const double d = widen_exact(a) - p;
    const bool gv = d > t.v;
    const bool gw = d > t.w;
    t.w = gv ? t.v : (gw ? d : t.w);
    t.v = gv ? d : t.v;
    t.j = gv ? col : t.j;
    t.a = gv ? a : t.a;
}

template <typename T> __device__ __forceinline__ void top2_merge(Top2<T>& t, const Top2<T>& o)
{
    const bool other = o.v > t.v || (o.v == t.v && o.j < t.j && o.j >= 0);
    const double lose_v = other ? t.v : o.v;
    const double keep_w = other ? o.w : t.w;
    t.w = lose_v > keep_w ? lose_v : keep_w;
    t.v = other ? o.v : t.v;
    t.j = other ? o.j : t.j;
    t.a = other ? o.a : t.a;
}

template <typename T> __device__ __forceinline__ Top2<T> top2_shfl_xor(const Top2<T>& t, int mask)
{
    Top2<T> o;
    o.v = __shfl_xor_sync(kFullMask, t.v, mask);
    o.w = __shfl_xor_sync(kFullMask, t.w, mask);
    o.j = __shfl_xor_sync(kFullMask, t.j, mask);
    o.a = __shfl_xor_sync(kFullMask, t.a, mask);
    return o;
}

// L2 read of a partial result another SM published (no stale L1 line).
template <typename T> __device__ __forceinline__ Top2<T> top2_load_cg(const Top2<T>* p)
{
    Top2<T> o;
    o.v = __ldcg(&p->v);
    o.w = __ldcg(&p->w);
    o.j = __ldcg(&p->j);
    o.a = __ldcg(&p->a);
    return o;
}

// All lanes end with the warp-wide result.
template <typename T> __device__ __forceinline__ void top2_warp_reduce(Top2<T>& t)
{
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1)
    {
        const Top2<T> o = top2_shfl_xor(t, m);
        top2_merge(t, o);
    }
}

// ---------------------------------------------------------------------------------------------
// Streaming loads of the benefit matrix: read once per bid, never written during a solve, so
// they take the non-coherent path and do not allocate in L1 (prices keep L1 to themselves).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 ld_stream4(const float* p)
{
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}

// Same, but the line is also first in line for eviction from L2: the cache-building sweep reads the
// whole matrix once and must not push the row caches it is writing (4 MB) out of the 126 MB L2.
__device__ __forceinline__ float4 ld_stream4_evict_first(const float* p)
{
    float4 r;
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p), "l"(pol));
    return r;
}

struct double4_t
{
    double x, y, z, w;
};

__device__ __forceinline__ double4_t ld_stream4(const double* p)
{
    double4_t r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(r.z), "=d"(r.w) : "l"(p + 2));
    return r;
}

__device__ __forceinline__ double4_t ld_stream4_evict_first(const double* p)
{
    double4_t r;
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;"
                 : "=d"(r.x), "=d"(r.y)
                 : "l"(p), "l"(pol));
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;"
                 : "=d"(r.z), "=d"(r.w)
                 : "l"(p + 2), "l"(pol));
    return r;
}

// exact widening of a coefficient of either type
__device__ __forceinline__ double widen_exact(double a) { return a; }

template <typename T> struct Vec4;
template <> struct Vec4<float>
{
    using type = float4;
};
template <> struct Vec4<double>
{
    using type = double4_t;
};

} // namespace b200lsap
