// Test program for the multi-GPU side of include/AuctionB200.h and include/b200lsap.h: no Python, no NCCL.
// Usage: multi_demo rows cols seed mode dev0 [dev1 ...]
//   mode sharded[_rowmajor]: one rows x cols problem column-sharded over the listed devices (an id may repeat: several
//        ranks on one GPU) through LSAP::AuctionShardedB200; prints "x y v" per edge and "objective <sum>", and checks
//        the edges against the single-GPU LSAP::AuctionB200 solve of the same matrix ("same_as_single 1")
//   mode batch: `rows` problems of 32 x 48 (seed, problem ids 0..) split over the devices through
//        b200lsap_multi_solve_batch_f32; prints one line of 32 columns per problem and "same_as_single 1"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "AuctionB200.h"
#include "b200lsap_gen.h"

template <bool RowMajor> struct Mat
{
    enum { IsRowMajor = RowMajor };
    size_t r, c;
    std::vector<float> d;
    Mat(size_t rows, size_t cols) : r(rows), c(cols), d(rows * cols) {}
    size_t rows() const { return r; }
    size_t cols() const { return c; }
    const float* data() const { return d.data(); }
    size_t outerStride() const { return RowMajor ? c : r; }
    float& operator()(size_t i, size_t j) { return RowMajor ? d[i * c + j] : d[i + j * r]; }
    const float& operator()(size_t i, size_t j) const { return RowMajor ? d[i * c + j] : d[i + j * r]; }
};

template <bool RowMajor> int run_sharded(size_t rows, size_t cols, unsigned long long seed, const std::vector<int>& devs)
{
    Mat<RowMajor> m(rows, cols);
    const uint64_t key = b200lsap_gen_key(seed, 0);
    for (size_t i = 0; i < rows; i++)
        for (size_t j = 0; j < cols; j++)
            m(i, j) = b200lsap_gen_value(key, i * cols + j);
    b200lsap_options o;
    b200lsap_options_default(&o);
    o.eps_start = 0.25; // square problems: the sharded solve's schedule, so that both sides play the same phases
    const auto sharded = LSAP::AuctionShardedB200<float>::solve(m, devs, &o);
    o.tail_rows = -1;
    o.narrow_rows = -1; // wide rounds only: the rounds of the sharded solve
    const auto single = LSAP::AuctionB200<float>::solve(m, &o);
    bool same = sharded.size() == single.size();
    double obj = 0;
    for (size_t k = 0; k < sharded.size(); k++)
    {
        same = same && sharded[k].x == single[k].x && sharded[k].y == single[k].y && sharded[k].v == single[k].v;
        obj += (double)sharded[k].v;
        std::printf("%zu %zu %.9g\n", sharded[k].x, sharded[k].y, (double)sharded[k].v);
    }
    std::printf("objective %.12f\nsame_as_single %d\n", obj, same ? 1 : 0);
    return same ? 0 : 3;
}

int run_batch(int64_t batch, unsigned long long seed, const std::vector<int>& devs)
{
    const int rows = 32, cols = 48;
    std::vector<float> a((size_t)batch * rows * cols);
    for (int64_t b = 0; b < batch; b++)
    {
        const uint64_t key = b200lsap_gen_key(seed, (uint64_t)b);
        for (int e = 0; e < rows * cols; e++)
            a[(size_t)b * rows * cols + e] = b200lsap_gen_value(key, (uint64_t)e);
    }
    std::vector<int32_t> multi((size_t)batch * rows, -1), single((size_t)batch * rows, -2);
    b200lsap_multi* m = nullptr;
    int rc = b200lsap_create_multi(&m, devs.data(), (int)devs.size());
    if (rc)
    {
        std::printf("create_multi: %s\n", b200lsap_status_string(rc));
        return 1;
    }
    b200lsap_stats st;
    rc = b200lsap_multi_solve_batch_f32(m, a.data(), batch, rows, cols, B200LSAP_ROW_MAJOR, nullptr, multi.data(), &st);
    if (rc)
    {
        std::printf("multi_solve_batch: %s\n", b200lsap_multi_last_error(m));
        return 1;
    }
    b200lsap_destroy_multi(m);
    b200lsap_handle* h = nullptr;
    if (b200lsap_create(&h, devs[0]) || b200lsap_solve_batch_f32(h, a.data(), batch, rows, cols, B200LSAP_ROW_MAJOR, nullptr,
                                                                 single.data(), nullptr))
        return 1;
    b200lsap_destroy(h);
    for (int64_t b = 0; b < batch; b++)
    {
        for (int i = 0; i < rows; i++)
            std::printf("%d%c", multi[(size_t)b * rows + i], i + 1 < rows ? ' ' : '\n');
    }
    const bool same = multi == single;
    std::printf("same_as_single %d\n", same ? 1 : 0);
    return same ? 0 : 3;
}

int main(int argc, char** argv)
{
    if (argc < 6)
    {
        std::printf("usage: multi_demo rows cols seed sharded|sharded_rowmajor|batch dev0 [dev1 ...]\n");
        return 2;
    }
    const size_t rows = (size_t)atoll(argv[1]), cols = (size_t)atoll(argv[2]);
    const unsigned long long seed = strtoull(argv[3], 0, 10);
    const std::string mode = argv[4];
    std::vector<int> devs;
    for (int k = 5; k < argc; k++)
        devs.push_back(atoi(argv[k]));
    try
    {
        if (mode == "batch")
            return run_batch((int64_t)rows, seed, devs);
        return mode == "sharded_rowmajor" ? run_sharded<true>(rows, cols, seed, devs) : run_sharded<false>(rows, cols, seed, devs);
    }
    catch (const std::exception& e)
    {
        std::printf("error: %s\n", e.what());
        return 1;
    }
}
