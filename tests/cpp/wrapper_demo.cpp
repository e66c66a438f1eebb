// Test program for include/AuctionB200.h.
//   default build: a minimal dense matrix class (the GPU box has no Eigen);
//   -DWITH_REFERENCE_EIGEN: the reference's own headers and Eigen types (compile check here).
// Usage: wrapper_demo rows cols seed [rowmajor] -> prints "x y v" per edge, then "objective <sum>".
#ifdef WITH_REFERENCE_EIGEN
#include "AuctionAlgorithm/Auction.h"
#include "AuctionAlgorithm/AuctionMT.h"
#endif
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "AuctionB200.h"
#include "b200lsap_gen.h"

template <typename S, bool RowMajor> struct MiniMatrix
{
    enum { IsRowMajor = RowMajor };
    size_t r, c;
    std::vector<S> d;
    MiniMatrix(size_t rows, size_t cols) : r(rows), c(cols), d(rows * cols) {}
    size_t rows() const { return r; }
    size_t cols() const { return c; }
    const S* data() const { return d.data(); }
    size_t outerStride() const { return RowMajor ? c : r; }
    S& operator()(size_t i, size_t j) { return RowMajor ? d[i * c + j] : d[i + j * r]; }
    const S& operator()(size_t i, size_t j) const { return RowMajor ? d[i * c + j] : d[i + j * r]; }
};

template <typename M> void fill(M& m, uint64_t seed)
{
    const uint64_t key = b200lsap_gen_key(seed, 0);
    for (size_t i = 0; i < (size_t)m.rows(); ++i)
        for (size_t j = 0; j < (size_t)m.cols(); ++j)
            m(i, j) = b200lsap_gen_value(key, i * m.cols() + j);
}

template <typename Edges> void print(const Edges& e)
{
    double obj = 0;
    for (const auto& edge : e)
    {
        printf("%zu %zu %.9g\n", edge.x, edge.y, (double)edge.v);
        obj += edge.v;
    }
    printf("objective %.12f\n", obj);
}

int main(int argc, char** argv)
{
    if (argc < 4)
        return 2;
    const size_t rows = atol(argv[1]), cols = atol(argv[2]);
    const uint64_t seed = atol(argv[3]);
    const bool rowmajor = argc > 4;
    try
    {
#ifdef WITH_REFERENCE_EIGEN
        Eigen::MatrixXd m(rows, cols);
        fill(m, seed);
        LSAP::Auction<double>::Edges solution = LSAP::AuctionB200<double>::solve(m);   // cf. LSAP.cpp:44
        solution = LSAP::AuctionMTB200<double>::solve(m, 2);                           // cf. LSAP.cpp:56
        Eigen::MatrixXf mf = m.cast<float>();
        LSAP::Auction<float, Eigen::MatrixXf>::Edges sf = LSAP::AuctionB200<float>::solve(mf);
        (void)sf;
        print(solution);
#else
        if (rowmajor)
        {
            MiniMatrix<float, true> m(rows, cols);
            fill(m, seed);
            print(LSAP::AuctionB200<float>::solve(m));
        }
        else
        {
            MiniMatrix<double, false> m(rows, cols);
            fill(m, seed);
            print(LSAP::AuctionMTB200<double>::solve(m, 2));
        }
#endif
    }
    catch (const std::exception& e)
    {
        printf("error %s\n", e.what());
        return 1;
    }
    return 0;
}
