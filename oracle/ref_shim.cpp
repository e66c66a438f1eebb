/*
 * TEST INFRASTRUCTURE — C entry points around the UNMODIFIED reference headers.
 *
 * Compiled by oracle/Makefile from the sources where they lie under $(REFERENCE_DIR)
 * (default /root/reference) into oracle/_ref/libref_auction.so.  Nothing of the reference is
 * copied: this file only instantiates LSAP::Auction / LSAP::AuctionMT (Auction.h:44,
 * AuctionMT.h:50) on Eigen matrices it fills from plain pointers, and flattens the
 * returned Edges (AuctionCommon.h:30-40).  Never built with -DDEBUG_AUCTION.
 *
 * The reference has no error path: rows > cols or an all-zero row never return
 * (SURVEY.md §0.5).  Callers keep to rows <= cols and strictly positive inputs.
 */
#include <thread>

#include "AuctionAlgorithm/Auction.h"
#include "AuctionAlgorithm/AuctionMT.h"

#include <cstdint>
#include <cstring>
#include <vector>

namespace
{

template <typename S> using Mat = Eigen::Matrix<S, -1, -1>;

template <typename S> Mat<S> to_eigen(const S* a, int64_t rows, int64_t cols, int64_t ld, int layout)
{
    Mat<S> m(rows, cols);
    for (int64_t j = 0; j < cols; j++)
        for (int64_t i = 0; i < rows; i++)
            m(i, j) = layout ? a[i * ld + j] : a[i + j * ld];
    return m;
}

template <typename S, typename Edges>
int64_t flatten(const Edges& e, int64_t* ex, int64_t* ey, S* ev)
{
    int64_t n = 0;
    for (const auto& edge : e)
    {
        ex[n] = (int64_t)edge.x;
        ey[n] = (int64_t)edge.y;
        ev[n] = edge.v;
        n++;
    }
    return n;
}

} // namespace

extern "C" {

/* returns the number of edges; seconds_out (nullable) = wall time of solve() alone, the
 * region the reference's own driver times (LSAP.cpp:44). */
int64_t ref_auction_f64(const double* a, int64_t rows, int64_t cols, int64_t ld, int layout, int64_t* ex,
                        int64_t* ey, double* ev, double* seconds_out)
{
    const Mat<double> m = to_eigen<double>(a, rows, cols, ld, layout);
    const auto t0 = std::chrono::steady_clock::now();
    const auto edges = LSAP::Auction<double>::solve(m);
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds_out)
        *seconds_out = std::chrono::duration<double>(t1 - t0).count();
    return flatten<double>(edges, ex, ey, ev);
}

int64_t ref_auction_f32(const float* a, int64_t rows, int64_t cols, int64_t ld, int layout, int64_t* ex,
                        int64_t* ey, float* ev, double* seconds_out)
{
    const Mat<float> m = to_eigen<float>(a, rows, cols, ld, layout);
    const auto t0 = std::chrono::steady_clock::now();
    const auto edges = LSAP::Auction<float, Eigen::MatrixXf>::solve(m);
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds_out)
        *seconds_out = std::chrono::duration<double>(t1 - t0).count();
    return flatten<float>(edges, ex, ey, ev);
}

int64_t ref_auctionmt_f64(const double* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                          int64_t n_threads, int64_t* ex, int64_t* ey, double* ev, double* seconds_out)
{
    const Mat<double> m = to_eigen<double>(a, rows, cols, ld, layout);
    const auto t0 = std::chrono::steady_clock::now();
    const auto edges = LSAP::AuctionMT<double>::solve(m, (size_t)n_threads);
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds_out)
        *seconds_out = std::chrono::duration<double>(t1 - t0).count();
    return flatten<double>(edges, ex, ey, ev);
}

/* fp32 data widened exactly to double, then Auction<double>: the oracle instantiation for
 * every fp32 config. */
int64_t ref_auction_f32_as_f64(const float* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                               int64_t* ex, int64_t* ey, double* ev, double* seconds_out)
{
    Mat<double> m(rows, cols);
    for (int64_t j = 0; j < cols; j++)
        for (int64_t i = 0; i < rows; i++)
            m(i, j) = (double)(layout ? a[i * ld + j] : a[i + j * ld]);
    const auto t0 = std::chrono::steady_clock::now();
    const auto edges = LSAP::Auction<double>::solve(m);
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds_out)
        *seconds_out = std::chrono::duration<double>(t1 - t0).count();
    return flatten<double>(edges, ex, ey, ev);
}

/* The reference has no batch API: Auction<double>::solve called once per problem, the problems
 * split over n_threads std::threads (solve() is re-entrant: locals only).  a: batch contiguous row-major
 * rows x cols fp32 problems, widened exactly to double per problem.  row_to_col[b*rows + i] = column of row i.
 * seconds_out: wall time of the whole threaded loop (matrix fill + solve per problem). */
int64_t ref_auction_batch_f32_as_f64(const float* a, int64_t batch, int rows, int cols, int n_threads,
                                     int32_t* row_to_col, double* seconds_out)
{
    if (n_threads < 1)
        n_threads = 1;
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < n_threads; t++)
        pool.emplace_back([=]() {
            Mat<double> m(rows, cols);
            for (int64_t b = batch * t / n_threads; b < batch * (t + 1) / n_threads; b++)
            {
                const float* p = a + b * (int64_t)rows * cols;
                for (int i = 0; i < rows; i++)
                    for (int j = 0; j < cols; j++)
                        m(i, j) = (double)p[i * cols + j];
                const auto edges = LSAP::Auction<double>::solve(m);
                for (const auto& e : edges)
                    row_to_col[b * rows + (int64_t)e.x] = (int32_t)e.y;
            }
        });
    for (auto& th : pool)
        th.join();
    if (seconds_out)
        *seconds_out = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return batch;
}

/* Auction<double, SparseMatrix<double, RowMajor>>::solve (the driver's second call, LSAP.cpp:36-38,51)
 * on a compressed row-major matrix given as Eigen's own three arrays. */
int64_t ref_auction_csr_f64(const double* values, const int32_t* row_ptr, const int32_t* col_idx, int64_t rows,
                            int64_t cols, int64_t* ex, int64_t* ey, double* ev, double* seconds_out)
{
    typedef Eigen::SparseMatrix<double, Eigen::RowMajor> Sp;
    const int64_t nnz = row_ptr[rows];
    const Sp m = Eigen::Map<const Sp>(rows, cols, nnz, row_ptr, col_idx, values);
    const auto t0 = std::chrono::steady_clock::now();
    const auto edges = LSAP::Auction<double, Sp>::solve(m);
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds_out)
        *seconds_out = std::chrono::duration<double>(t1 - t0).count();
    return flatten<double>(edges, ex, ey, ev);
}

} // extern "C"
