/*
 * AuctionB200.h — header-only C++ mirror of the reference's solver interface over the C-ABI
 * (b200lsap.h).  Same names, argument meaning and result types as
 *
 *   LSAP::Auction<Scalar, MatrixType>::solve(const MatrixType&)            Auction.h:19-44
 *   LSAP::AuctionMT<Scalar, MatrixType>::solve(const MatrixType&, size_t)  AuctionMT.h:24-50
 *   LSAP::Auction<Scalar, Eigen::SparseMatrix<Scalar, Eigen::RowMajor>>::solve   Auction.h:166-198
 *   LSAP::AuctionCommon<Scalar>::Edge / Edges                             AuctionCommon.h:30-40
 *
 * so a caller such as the reference's driver (LSAP.cpp:41-45,56) only changes the class name:
 *
 *   Auction<double>::Edges solution = LSAP::AuctionB200<double>::solve(m);
 *   solution = LSAP::AuctionMTB200<double>::solve(m, 2);
 *
 * MatrixType is any dense matrix with rows(), cols(), data(), outerStride() and operator()(i,j)
 * plus a compile-time IsRowMajor (every Eigen::Matrix<Scalar,-1,-1[,RowMajor]> qualifies; Eigen
 * itself is not required by this header).  When AuctionCommon.h has been included before this
 * header the reference's own Edge type is used, otherwise an identical one is defined.
 *
 * Differences from the reference, all on inputs where the reference never returns:
 * rows > cols, a row without a non-zero coefficient, or an expired round watchdog throw
 * std::runtime_error.  Edges come back ordered by row (e.x ascending).
 */
#ifndef AUCTION_B200_H_
#define AUCTION_B200_H_

#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "b200lsap.h"

namespace LSAP
{

#ifdef AUCTIONCOMMON_H_
template <typename Scalar> struct B200Types
{
    typedef typename AuctionCommon<Scalar>::Edge Edge;
    typedef typename AuctionCommon<Scalar>::Edges Edges;
};
#else
template <typename Scalar> struct B200Types
{
    struct Edge
    {
        size_t x;
        size_t y;
        Scalar v;
    };
    typedef std::vector<Edge> Edges;
};
#endif

namespace b200detail
{

/* one handle per host thread: handles serve one thread at a time (b200lsap.h) */
inline b200lsap_handle* thread_handle()
{
    struct Holder
    {
        b200lsap_handle* h = nullptr;
        ~Holder() { b200lsap_destroy(h); }
    };
    static thread_local Holder holder;
    if (!holder.h)
    {
        const int rc = b200lsap_create(&holder.h, -1);
        if (rc != B200LSAP_OK)
            throw std::runtime_error(std::string("b200lsap_create: ") + b200lsap_status_string(rc));
    }
    return holder.h;
}

inline int solve_raw(b200lsap_handle* h, const float* a, int64_t r, int64_t c, int64_t ld, int layout,
                     const b200lsap_options* o, int64_t* out, b200lsap_stats* st)
{
    return b200lsap_solve_f32(h, a, r, c, ld, layout, o, out, nullptr, st);
}
inline int solve_raw(b200lsap_handle* h, const double* a, int64_t r, int64_t c, int64_t ld, int layout,
                     const b200lsap_options* o, int64_t* out, b200lsap_stats* st)
{
    return b200lsap_solve_f64(h, a, r, c, ld, layout, o, out, nullptr, st);
}

inline int solve_csr_raw(b200lsap_handle* h, const float* v, const int32_t* rp, const int32_t* ci, int64_t r, int64_t c,
                         int64_t nnz, const b200lsap_options* o, int64_t* out, b200lsap_stats* st)
{
    return b200lsap_solve_csr_f32(h, v, rp, ci, r, c, nnz, o, out, nullptr, st);
}
inline int solve_csr_raw(b200lsap_handle* h, const double* v, const int32_t* rp, const int32_t* ci, int64_t r, int64_t c,
                         int64_t nnz, const b200lsap_options* o, int64_t* out, b200lsap_stats* st)
{
    return b200lsap_solve_csr_f64(h, v, rp, ci, r, c, nnz, o, out, nullptr, st);
}

/* a compressed sparse matrix type (Eigen::SparseMatrix): it has innerIndexPtr() */
template <typename M> struct is_sparse
{
    template <typename U> static char test(decltype(((const U*)0)->innerIndexPtr())*);
    template <typename U> static long test(...);
    enum
    {
        value = sizeof(test<M>(0)) == sizeof(char)
    };
};
template <bool B> struct bool_tag
{
};

} // namespace b200detail

template <typename Scalar = double, typename MatrixType = void> class AuctionB200
{
public:
    typedef typename B200Types<Scalar>::Edge Edge;
    typedef typename B200Types<Scalar>::Edges Edges;

    /* Replaces Auction<Scalar,MatrixType>::solve (Auction.h:44) for a dense matrix and for
     * Eigen::SparseMatrix<Scalar, Eigen::RowMajor> (findForward on stored entries, Auction.h:166-198;
     * the driver's Auction<double, SparseMatrix<double, RowMajor>>::solve(s), LSAP.cpp:51).
     * opts / stats optional. */
    template <typename M>
    static const Edges solve(const M& a, const b200lsap_options* opts = nullptr, b200lsap_stats* stats = nullptr)
    {
        return solve_impl(a, opts, stats, b200detail::bool_tag<b200detail::is_sparse<M>::value>());
    }

private:
    template <typename M>
    static const Edges solve_impl(const M& a, const b200lsap_options* opts, b200lsap_stats* stats, b200detail::bool_tag<false>)
    {
        const int64_t rows = (int64_t)a.rows(), cols = (int64_t)a.cols();
        Edges E;
        if (rows == 0 || cols == 0)
            return E;
        std::vector<int64_t> row_to_col((size_t)rows, -1);
        const int layout = M::IsRowMajor ? B200LSAP_ROW_MAJOR : B200LSAP_COL_MAJOR;
        b200lsap_handle* h = b200detail::thread_handle();
        const Scalar* data = a.data();
        const int rc = b200detail::solve_raw(h, data, rows, cols, (int64_t)a.outerStride(), layout, opts,
                                             row_to_col.data(), stats);
        if (rc != B200LSAP_OK)
            throw std::runtime_error(std::string("AuctionB200::solve: ") + b200lsap_last_error(h));
        E.resize((size_t)rows);
        for (int64_t i = 0; i < rows; ++i)
        {
            E[(size_t)i].x = (size_t)i;
            E[(size_t)i].y = (size_t)row_to_col[(size_t)i];
            E[(size_t)i].v = a(i, row_to_col[(size_t)i]); /* from the caller's matrix: Auction.h:273,283 */
        }
        return E;
    }

    /* compressed row-major sparse matrix (the reference asserts isCompressed(), Auction.h:170) */
    template <typename M>
    static const Edges solve_impl(const M& a, const b200lsap_options* opts, b200lsap_stats* stats, b200detail::bool_tag<true>)
    {
        static_assert(M::IsRowMajor, "the reference's sparse path is row-major (Auction.h:166)");
        static_assert(sizeof(*((const M*)0)->innerIndexPtr()) == 4, "StorageIndex must be a 32-bit integer");
        const int64_t rows = (int64_t)a.rows(), cols = (int64_t)a.cols();
        Edges E;
        if (rows == 0 || cols == 0)
            return E;
        if (!a.isCompressed())
            throw std::runtime_error("AuctionB200::solve: the sparse matrix must be compressed (makeCompressed())");
        std::vector<int64_t> row_to_col((size_t)rows, -1);
        b200lsap_handle* h = b200detail::thread_handle();
        const Scalar* values = a.valuePtr();
        const int rc = b200detail::solve_csr_raw(h, values, (const int32_t*)a.outerIndexPtr(),
                                                 (const int32_t*)a.innerIndexPtr(), rows, cols, (int64_t)a.nonZeros(),
                                                 opts, row_to_col.data(), stats);
        if (rc != B200LSAP_OK)
            throw std::runtime_error(std::string("AuctionB200::solve: ") + b200lsap_last_error(h));
        E.resize((size_t)rows);
        for (int64_t i = 0; i < rows; ++i)
        {
            E[(size_t)i].x = (size_t)i;
            E[(size_t)i].y = (size_t)row_to_col[(size_t)i];
            E[(size_t)i].v = a.coeff(i, row_to_col[(size_t)i]); /* the stored value: Auction.h:273,283 */
        }
        return E;
    }
};

/* AuctionMT<Scalar,MatrixType>::solve(m, noOfThreads) (AuctionMT.h:50): noOfThreads is accepted
 * and ignored — the column-interval split it sized (AuctionMT.h:76-77) is the GPU grid's. */
template <typename Scalar = double, typename MatrixType = void> class AuctionMTB200
{
public:
    typedef typename B200Types<Scalar>::Edge Edge;
    typedef typename B200Types<Scalar>::Edges Edges;

    template <typename M> static Edges solve(const M& m, const size_t noOfThreads = 2)
    {
        (void)noOfThreads;
        return AuctionB200<Scalar>::solve(m);
    }
};

/* The same solve with the matrix column-sharded over several GPUs of this machine (AuctionMT's column intervals,
 * AuctionMT.h:76-77, with a GPU per interval and one in-kernel exchange per round; b200lsap_multi_solve_f32).
 * fp32 matrices.  devices: CUDA device ids, one rank each (an id may repeat).  The b200lsap_multi is built on first
 * use and kept per host thread for the same device list.
 *
 *   Auction<float, Eigen::MatrixXf>::Edges e = LSAP::AuctionShardedB200<float>::solve(m, {0, 1, 2, 3});
 */
template <typename Scalar = float, typename MatrixType = void> class AuctionShardedB200
{
public:
    typedef typename B200Types<Scalar>::Edge Edge;
    typedef typename B200Types<Scalar>::Edges Edges;

    template <typename M>
    static Edges solve(const M& a, const std::vector<int>& devices, const b200lsap_options* opts = nullptr,
                       b200lsap_stats* stats = nullptr)
    {
        static_assert(sizeof(Scalar) == 4, "the column-sharded solve takes fp32 matrices");
        struct Holder
        {
            b200lsap_multi* m = nullptr;
            std::vector<int> devices;
            ~Holder() { b200lsap_destroy_multi(m); }
        };
        static thread_local Holder holder;
        if (!holder.m || holder.devices != devices)
        {
            b200lsap_destroy_multi(holder.m);
            holder.m = nullptr;
            const int rc = b200lsap_create_multi(&holder.m, devices.data(), (int)devices.size());
            if (rc != B200LSAP_OK)
                throw std::runtime_error(std::string("b200lsap_create_multi: ") + b200lsap_status_string(rc));
            holder.devices = devices;
        }
        const int64_t rows = (int64_t)a.rows(), cols = (int64_t)a.cols();
        Edges E;
        if (rows == 0 || cols == 0)
            return E;
        std::vector<int64_t> row_to_col((size_t)rows, -1);
        const int layout = M::IsRowMajor ? B200LSAP_ROW_MAJOR : B200LSAP_COL_MAJOR;
        const int rc = b200lsap_multi_solve_f32(holder.m, a.data(), rows, cols, (int64_t)a.outerStride(), layout, opts,
                                                row_to_col.data(), nullptr, stats);
        if (rc != B200LSAP_OK)
            throw std::runtime_error(std::string("AuctionShardedB200::solve: ") + b200lsap_multi_last_error(holder.m));
        E.resize((size_t)rows);
        for (int64_t i = 0; i < rows; ++i)
        {
            E[(size_t)i].x = (size_t)i;
            E[(size_t)i].y = (size_t)row_to_col[(size_t)i];
            E[(size_t)i].v = a(i, row_to_col[(size_t)i]); /* from the caller's matrix: Auction.h:273,283 */
        }
        return E;
    }
};

} // namespace LSAP

#endif /* AUCTION_B200_H_ */
