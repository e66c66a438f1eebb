// Test program for include/AuctionB200.h.
//   default build: a minimal dense matrix class (the GPU box has no Eigen);
//   -DWITH_REFERENCE_EIGEN: the reference's own headers and Eigen types (compile check here).
// Usage: wrapper_demo rows cols seed [rowmajor|sparse] -> prints "x y v" per edge, then "objective <sum>".
//   sparse: a compressed row-major matrix keeping entry (i, j) when j == i or (i * 7 + j * 13) % 4 != 0.
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

// what AuctionB200 needs of Eigen::SparseMatrix<S, RowMajor> (compressed)
template <typename S> struct MiniSparse
{
    enum { IsRowMajor = 1 };
    size_t r, c;
    std::vector<int> ptr, idx;
    std::vector<S> val;
    template <typename D> explicit MiniSparse(const D& d) : r(d.rows()), c(d.cols()), ptr(1, 0)
    {
        for (size_t i = 0; i < r; ++i)
        {
            for (size_t j = 0; j < c; ++j)
                if (j == i || (i * 7 + j * 13) % 4 != 0)
                {
                    idx.push_back((int)j);
                    val.push_back(d(i, j));
                }
            ptr.push_back((int)idx.size());
        }
    }
    size_t rows() const { return r; }
    size_t cols() const { return c; }
    bool isCompressed() const { return true; }
    size_t nonZeros() const { return val.size(); }
    const S* valuePtr() const { return val.data(); }
    const int* innerIndexPtr() const { return idx.data(); }
    const int* outerIndexPtr() const { return ptr.data(); }
    S coeff(size_t i, size_t j) const
    {
        for (int k = ptr[i]; k < ptr[i + 1]; ++k)
            if (idx[k] == (int)j)
                return val[k];
        return S(0);
    }
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
    const bool sparse = argc > 4 && argv[4][0] == 's';
    const bool rowmajor = argc > 4 && !sparse;
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
        Eigen::SparseMatrix<double, Eigen::RowMajor> sp(rows, cols);
        sp = m.sparseView();
        sp.makeCompressed();
        LSAP::Auction<double>::Edges ss = LSAP::AuctionB200<double>::solve(sp);         // cf. LSAP.cpp:51
        (void)ss;
        print(solution);
#else
        if (sparse)
        {
            MiniMatrix<double, true> m(rows, cols);
            fill(m, seed);
            const MiniSparse<double> sp(m);
            print(LSAP::AuctionB200<double>::solve(sp));
        }
        else if (rowmajor)
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
