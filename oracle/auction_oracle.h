/*
 * TEST INFRASTRUCTURE — C interface of the CPU oracle (see auction_oracle_impl.h for what
 * it restates, with reference file:line).  Parity status: PINNED — the oracle is checked
 * against the unmodified reference compiled into oracle/_ref (tests/test_oracle.py) and
 * against golden vectors the reference produced (tests/golden/, made by make_golden.py).
 *
 * Not part of the product: only tests/, __graft_entry__.smoke() and bench.py's CPU legs
 * may load liboracle.so.
 */
#ifndef AUCTION_ORACLE_H_
#define AUCTION_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORACLE_EPS_MULT 1e-5 /* internal.h:16 */
#define ORACLE_INF 1e6       /* internal.h:17 */

#define ORACLE_COL_MAJOR 0 /* Eigen default */
#define ORACLE_ROW_MAJOR 1

enum
{
    ORC_OK = 0,
    ORC_BUDGET = 1,    /* max_bids ran out before termination */
    ORC_INFEASIBLE = 2 /* the reference would spin forever (rows > cols, empty row) */
};

typedef struct
{
    int64_t sweeps;         /* forward() calls, Auction.h:210 */
    int64_t bids;           /* rows that placed a bid, Auction.h:250 */
    int64_t evictions;      /* edges overwritten, Auction.h:267-277 */
    int64_t scan_steps;     /* iterations of the edge scan, Auction.h:267 */
    int64_t lambda_rejects; /* else-branch Auction.h:291-294 */
    int64_t reverse_calls;  /* reverse() calls, Auction.h:346 */
    int64_t reverse_work;   /* columns reverse() actually bid for */
    int64_t status;
} oracle_stats;

/* Auction<double, MatrixXd>::solve.  a(i,j) = a[i + j*ld] (layout 0) or a[i*ld + j] (layout 1).
 * max_bids < 0: unlimited.  Edges come back in the reference's order (first win of a column).
 * edge_* must hold min(rows, cols) entries (cols when rows > cols).  prices_out / profits_out
 * may be NULL. */
int oracle_auction_f64(const double* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                       int64_t max_bids, int64_t* edge_x, int64_t* edge_y, double* edge_v,
                       int64_t* n_edges, double* prices_out, double* profits_out, oracle_stats* stats);

/* Auction<float, MatrixXf>::solve: every operation in float.  Livelocks on large or square
 * inputs — always pass a bid budget. */
int oracle_auction_f32(const float* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                       int64_t max_bids, int64_t* edge_x, int64_t* edge_y, float* edge_v,
                       int64_t* n_edges, float* prices_out, float* profits_out, oracle_stats* stats);

/* Auction<Scalar, Eigen::SparseMatrix<Scalar, Eigen::RowMajor>>::solve (Auction.h:166-198 under the
 * same controller) on a compressed row-major matrix: values / col_idx / row_ptr are Eigen's
 * valuePtr() / innerIndexPtr() / outerIndexPtr(), columns ascending inside a row.  Every stored
 * entry is an edge. */
int oracle_auction_csr_f64(const double* values, const int32_t* row_ptr, const int32_t* col_idx,
                           int64_t rows, int64_t cols, int64_t max_bids, int64_t* edge_x, int64_t* edge_y,
                           double* edge_v, int64_t* n_edges, double* prices_out, double* profits_out,
                           oracle_stats* stats);
int oracle_auction_csr_f32(const float* values, const int32_t* row_ptr, const int32_t* col_idx,
                           int64_t rows, int64_t cols, int64_t max_bids, int64_t* edge_x, int64_t* edge_y,
                           float* edge_v, int64_t* n_edges, float* prices_out, float* profits_out,
                           oracle_stats* stats);

/* Synthetic inputs (include/b200lsap_gen.h): fills a rows x cols matrix with leading
 * dimension ld in the given layout. */
void oracle_generate_f32(float* a, int64_t rows, int64_t cols, int64_t ld, int layout, uint64_t seed,
                         uint64_t problem);
void oracle_generate_f64(double* a, int64_t rows, int64_t cols, int64_t ld, int layout, uint64_t seed,
                         uint64_t problem);

/* The reference has no batch API: this is Auction<float>::solve or Auction<double>::solve (on
 * the widened fp32 data) looped over `batch` contiguous row-major rows x cols problems, split
 * over n_threads pthreads.  row_to_col[b*rows + i] = column of row i.  use_double selects the
 * arithmetic.  Returns the number of problems that did not terminate within the bid budget. */
int64_t oracle_auction_batch_f32(const float* a, int64_t batch, int rows, int cols, int use_double,
                                 int n_threads, int32_t* row_to_col);

#ifdef __cplusplus
}
#endif
#endif
