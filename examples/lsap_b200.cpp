// lsap_b200 — the reference's driver (LSAP.cpp:21-63) against the B200 path: the same four solves
// (dense and compressed row-major sparse, "single threaded" and "2 threads" class names), the same two
// output lines per solve ("command  <...> took N us" / "objective function value: X", internal.h:26-34).
//
//   g++ -std=c++11 -O2 -I include examples/lsap_b200.cpp -o lsap_b200 \
//       -L auctionalgorithm_b200 -lb200lsap -Wl,-rpath,$PWD/auctionalgorithm_b200
//   ./lsap_b200 [rows cols [seed [repeats]]]           default 5000 5000 (LSAP.cpp:23)
//
// With -DWITH_REFERENCE_EIGEN -I<reference> -I<reference>/external the matrices are Eigen's and the
// reference's own solvers run beside ours on the same input (A/B lines); without it two minimal
// containers with the members AuctionB200.h needs stand in for Eigen (the GPU box has none).
// Input recipe of LSAP.cpp:26-33: uniform coefficients in (-1, 1), shifted by +1, divided by the
// largest, i.e. values in (0, 1] with the maximum exactly 1.
#ifdef WITH_REFERENCE_EIGEN
#include "AuctionAlgorithm/Auction.h"
#include "AuctionAlgorithm/AuctionMT.h"
#endif
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <random>
#include <vector>

#include "AuctionB200.h"

using namespace LSAP;
typedef double Scalar;

#ifdef WITH_REFERENCE_EIGEN
typedef Eigen::Matrix<Scalar, -1, -1> Dense;
typedef Eigen::SparseMatrix<Scalar, Eigen::RowMajor> Sparse;
static Sparse to_sparse(const Dense& m)
{
    Sparse s(m.rows(), m.cols());
    s = m.sparseView();
    s.makeCompressed();
    return s;
}
#else
struct Dense // column-major like Eigen's default
{
    enum { IsRowMajor = 0 };
    size_t r, c;
    std::vector<Scalar> d;
    Dense(size_t rows, size_t cols) : r(rows), c(cols), d(rows * cols) {}
    size_t rows() const { return r; }
    size_t cols() const { return c; }
    const Scalar* data() const { return d.data(); }
    size_t outerStride() const { return r; }
    Scalar& operator()(size_t i, size_t j) { return d[i + j * r]; }
    const Scalar& operator()(size_t i, size_t j) const { return d[i + j * r]; }
};
struct Sparse // compressed row-major
{
    enum { IsRowMajor = 1 };
    size_t r, c;
    std::vector<int> ptr, idx;
    std::vector<Scalar> val;
    size_t rows() const { return r; }
    size_t cols() const { return c; }
    bool isCompressed() const { return true; }
    size_t nonZeros() const { return val.size(); }
    const Scalar* valuePtr() const { return val.data(); }
    const int* innerIndexPtr() const { return idx.data(); }
    const int* outerIndexPtr() const { return ptr.data(); }
    Scalar coeff(size_t i, size_t j) const
    {
        int lo = ptr[i], hi = ptr[i + 1];
        while (lo < hi)
        {
            const int mid = lo + (hi - lo) / 2;
            if (idx[mid] < (int)j)
                lo = mid + 1;
            else
                hi = mid;
        }
        return (lo < ptr[i + 1] && idx[lo] == (int)j) ? val[lo] : Scalar(0);
    }
};
static Sparse to_sparse(const Dense& m) // what sparseView() keeps: the non-zero coefficients
{
    Sparse s;
    s.r = m.rows();
    s.c = m.cols();
    s.ptr.push_back(0);
    for (size_t i = 0; i < s.r; ++i)
    {
        for (size_t j = 0; j < s.c; ++j)
            if (m(i, j) != 0)
            {
                s.idx.push_back((int)j);
                s.val.push_back(m(i, j));
            }
        s.ptr.push_back((int)s.idx.size());
    }
    return s;
}
#endif

template <typename Edges> static Scalar objFuncValue(const Edges& edges)
{
    Scalar val = 0.;
    for (const auto& e : edges)
        val += e.v;
    return val;
}

template <typename F> static void measure(const char* what, F f)
{
    std::cout << "command  <" << what << "> took " << std::flush;
    const auto t0 = std::chrono::high_resolution_clock::now();
    const Scalar obj = f();
    const std::chrono::nanoseconds ns = std::chrono::high_resolution_clock::now() - t0;
    std::cout << (ns.count() / 1000.) << " us " << std::endl;
    std::cout.precision(12);
    std::cout << "objective function value: " << obj << std::endl;
    std::cout.precision(6);
}

int main(int argc, char** argv)
{
    const size_t rows = argc > 2 ? atol(argv[1]) : 5000, cols = argc > 2 ? atol(argv[2]) : 5000;
    const unsigned seed = argc > 3 ? atol(argv[3]) : 1;
    const int repeats = argc > 4 ? atoi(argv[4]) : 1;
    if (rows < 1 || rows > cols)
    {
        std::cerr << "need 1 <= rows <= cols (LSAP.cpp:25)" << std::endl;
        return 2;
    }
    Dense m(rows, cols);
    {
        std::mt19937_64 gen(seed);
        std::uniform_real_distribution<Scalar> uni(-1., 1.);
        Scalar mx = 0;
        for (size_t j = 0; j < cols; ++j)
            for (size_t i = 0; i < rows; ++i)
            {
                m(i, j) = uni(gen) + 1.;
                mx = m(i, j) > mx ? m(i, j) : mx;
            }
        for (size_t j = 0; j < cols; ++j)
            for (size_t i = 0; i < rows; ++i)
                m(i, j) /= mx;
    }
    const Sparse s = to_sparse(m);
    try
    {
        for (int r = 0; r < repeats; ++r) // the first solve of a process also creates the CUDA context
        {
            measure("solution = AuctionB200<double>::solve(m)", [&] { return objFuncValue(AuctionB200<double>::solve(m)); });
            measure("solution = AuctionB200<double>::solve(s)", [&] { return objFuncValue(AuctionB200<double>::solve(s)); });
            measure("solution = AuctionMTB200<double>::solve(m, 2)", [&] { return objFuncValue(AuctionMTB200<double>::solve(m, 2)); });
            measure("solution = AuctionMTB200<double>::solve(s, 2)", [&] { return objFuncValue(AuctionMTB200<double>::solve(s, 2)); });
        }
#ifdef WITH_REFERENCE_EIGEN
        if (getenv("LSAP_B200_RUN_REFERENCE"))
        {
            measure("solution = Auction<double>::solve(m)", [&] { return objFuncValue(Auction<double>::solve(m)); });
            measure("solution = Auction<double, SparseMatrix<double, RowMajor> >::solve(s)",
                    [&] { return objFuncValue(Auction<double, Sparse>::solve(s)); });
        }
#endif
    }
    catch (const std::exception& e)
    {
        std::cout << "error " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
