/*
 * b200lsap.h — C-ABI of the B200-native auction solver (libb200lsap.so).
 *
 * This is the drop-in boundary for the reference's dense solve path.  Every entry point
 * names the reference interface it replaces, relative to /root/reference/:
 *
 *   b200lsap_solve_f64 / _f32      LSAP::Auction<Scalar,MatrixType>::solve(const MatrixType&)
 *                                  AuctionAlgorithm/Auction.h:44-109 (controller),
 *                                  :124-152 findForward, :210-306 forward; and
 *                                  LSAP::AuctionMT<Scalar,MatrixType>::solve(m, noOfThreads)
 *                                  AuctionAlgorithm/AuctionMT.h:50-171, :335-451 parallelForward,
 *                                  AuctionAlgorithm/AuctionCommon.h:86-129 forwardGS
 *   b200lsap_solve_batch_f32       the same solve() called once per problem by the caller
 *                                  (the reference has no batch API; LSAP.cpp:44 is one call)
 *   b200lsap_*_device_*            same operations on matrices already resident in HBM
 *   b200lsap_options.eps_multiplier  __AUCTION_EPSILON_MULTIPLIER, internal.h:16
 *   B200LSAP_NO_SECOND_BEST          __AUCTION_INF, internal.h:17
 *   b200lsap_shard_*               AuctionMT's column-interval decomposition
 *                                  (AuctionMT.h:76-77, AuctionCommon.h:302-320) and the top-2
 *                                  merge of thread 0 (AuctionMT.h:377-398), across GPUs
 *
 * Plain C types only.  The caller owns every host buffer; a handle owns its device buffers
 * and stream, is reusable, and serves one host thread at a time.  Nothing here falls back
 * to the CPU: without a CUDA device b200lsap_create fails.
 *
 * Semantics kept from the reference: a(i,j) == 0 means "no edge" (Auction.h:134); ties in a
 * row go to the lowest column (strict '>' at Auction.h:136); the second-best value may equal
 * the best (Auction.h:147-148); eps = eps_multiplier / cols (Auction.h:55); bid = a - w + eps
 * evaluated in double (Auction.h:250).  Where the reference never returns (rows > cols, a row
 * without an edge) these functions return an error status instead.
 */
#ifndef B200LSAP_H_
#define B200LSAP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200LSAP_VERSION 103

#define B200LSAP_COL_MAJOR 0 /* Eigen::Matrix default: a(i,j) = a[i + j*ld] */
#define B200LSAP_ROW_MAJOR 1 /* a(i,j) = a[i*ld + j] */

#define B200LSAP_NO_SECOND_BEST 1e6 /* internal.h:17 */

typedef enum
{
    B200LSAP_OK = 0,
    B200LSAP_ERR_SHAPE = 1,      /* rows > cols, or a non-positive dimension, or ld too small */
    B200LSAP_ERR_INFEASIBLE = 2, /* some row has no non-zero coefficient */
    B200LSAP_ERR_MAX_ROUNDS = 3, /* watchdog: max_rounds auction rounds without termination */
    B200LSAP_ERR_CUDA = 4,       /* see b200lsap_last_error */
    B200LSAP_ERR_ARG = 5,        /* null pointer, bad option */
    B200LSAP_ERR_NO_DEVICE = 6
} b200lsap_status;

typedef struct b200lsap_handle b200lsap_handle;

typedef struct
{
    double eps_multiplier; /* final eps = eps_multiplier / cols; default 1e-5 (internal.h:16) */
    int32_t eps_scaling;   /* -1 auto (on iff rows == cols), 0 off (the reference's fixed eps), 1 on, 2 on with the
                              asymmetric repair after the last phase only.  On with rows < cols (persistent driver
                              only): a forward auction with scaled eps may end a phase with an unassigned object
                              priced above an assigned one, which breaks the post-condition of
                              AuctionCommon.h:347-354 — what the reference's reverse auction and lambda are for
                              (Auction.h:346-469).  It is restored at the end of every phase (1) or of the last one
                              (2) by raising every price below the highest unassigned price to it (prices stay
                              monotone), unassigning the rows whose object was raised and letting them bid again at
                              that phase's eps (stats.warm_fallback = number of such repairs).  Measured: 1 pays
                              where the fixed eps crawls (rank-one 2048 x 2052: 2.87 M rounds -> 0.11 M) and costs
                              ~3x the rounds where it does not (uniform noise); 2 costs ~5 % and gains nothing. */
    double eps_start;      /* first eps of the scaling schedule; <= 0: default (an eighth of the mean spread
                              of a row's best ~36 values, measured on 128 sampled rows, at most 0.25).
                              A default first phase that runs over its work budget
                              restarts from a larger eps (stats.tail_runs).  Only the LAST phase, at
                              eps_multiplier / cols, decides the result. */
    double eps_factor;     /* eps is divided by this between phases; <= 1: default (7) */
    int64_t max_rounds;    /* watchdog; <= 0: default */
    int32_t driver;        /* 0 persistent cooperative kernel (default); 1 one launch per phase
                              of a round, driven from the host (profiling and cross-checks) */
    int32_t tail_rows;     /* a round with at most this many bidders is finished by the Gauss-Seidel
                              tail on one SM, one bid at a time; 0: default (1), -1: never, at most
                              256.  Persistent driver only. */
    int32_t narrow_rows;   /* a round with at most this many bidders (and at most SMs - 1) runs as a
                              narrow round: one CTA per bidder row, a controller CTA instead of grid
                              syncs; 0: default (SMs - 1), -1: never.  Persistent driver only. */
    int32_t row_cache;     /* narrow rounds answer from per-row caches of the best columns instead of
                              scanning rows (persistent driver; bit-identical bids): 0 default (on),
                              -1 off */
    int32_t batch_bid_cap; /* batched mode: a problem that has placed this many bids is handed from its two
                              lanes to its whole warp, which continues the same Gauss-Seidel sweep ~5x faster
                              per bid (same bids, same result; it only bounds how long a price war can hold
                              a launch).  0: default (96), -1: never */
    const double* prices_in; /* warm start of the dense solve (b200lsap_solve_f32/_f64 and their _device variants): cols
                              object prices to start from instead of zeros (Auction.h:58) — e.g. the prices_out of the
                              previous frame of a tracker.  Host memory for the host entries, device memory for the
                              _device entries.  Non-negative, finite.  Every row still starts unassigned, so the result
                              satisfies eps-complementary slackness at eps_multiplier / cols exactly as a cold solve
                              does (same assignment whenever the optimum is unique by more than rows * eps).  The eps
                              schedule is kept (its first eps is sized at the imported prices): measured, a frame whose
                              coefficients moved a little takes 10-30 % fewer rounds than a cold solve; eps_scaling = 0
                              (one phase at the final eps) only pays when the prices are already this matrix's
                              equilibrium.  rows < cols: the post-condition "no unassigned column is priced above an
                              assigned one" (AuctionCommon.h:347-354) is restored afterwards as for eps scaling on
                              asymmetric problems (eps_scaling above; stats.warm_fallback).  NULL: cold start. */
} b200lsap_options;

typedef struct
{
    int64_t rounds;    /* synchronous auction rounds executed */
    int64_t bids;      /* sum over rounds of bidding rows */
    int64_t phases;    /* eps phases (1 without scaling) */
    int64_t bid_bytes; /* algorithmic bytes of the bidding phase: bids * cols * sizeof(element) */
    double eps_final;
    double ms_upload;  /* host->device copy (+ transpose when the input is column-major) */
    double ms_solve;   /* auction kernels only, CUDA events on the handle's stream */
    double ms_download;
    int64_t launches;  /* kernels launched by this call */
    int64_t tail_bids; /* with row caches: bids of one-bidder rounds (k_narrow's single-warp chain; included
                          in bids and in rounds).  Row-scanning paths: bids placed by the Gauss-Seidel tail
                          (included in bids, not in rounds) */
    int64_t tail_runs; /* row-scanning paths: times the tail took over.  With row caches: times the
                          data-sized schedule was restarted from a larger eps (first phase over budget) */
    int64_t narrow_rounds; /* rounds run as narrow rounds (included in rounds) */
    int64_t cache_hits;    /* narrow-round bids answered from the row cache alone */
    int64_t cache_misses;  /* narrow-round bids that re-scanned their row (and rebuilt its cache) */
    double eps_first;      /* eps of the first phase */
    int64_t batch_takeovers; /* batched mode: problems that reached batch_bid_cap and were finished by a warp */
    int64_t batch_max_bids;  /* batched mode: largest bid count of a problem (filtered kernel) */
    int64_t warm_fallback;   /* rows < cols with eps scaling or a warm start: floor-raise repairs of the price
                                post-condition (AuctionCommon.h:347-354) after the last phase; 0 = it held at once */
} b200lsap_stats;

/* device < 0: the current CUDA device.  Creates the handle's own non-blocking stream. */
int b200lsap_create(b200lsap_handle** handle, int device);
int b200lsap_destroy(b200lsap_handle* handle);
/* Run the handle's work on a caller-owned cudaStream_t (e.g. torch's current stream). */
int b200lsap_set_stream(b200lsap_handle* handle, void* cuda_stream);
const char* b200lsap_last_error(const b200lsap_handle* handle);
const char* b200lsap_status_string(int status);
void b200lsap_options_default(b200lsap_options* opts);

/*
 * Replaces Auction<Scalar>::solve / AuctionMT<Scalar>::solve on a dense host matrix
 * (Auction.h:44, AuctionMT.h:50).  a: rows x cols, leading dimension ld, in `layout`.
 * row_to_col[i] receives the column assigned to row i (always >= 0 on success).  The caller
 * rebuilds Edge{x=i, y=row_to_col[i], v=a(i,y)} from its own matrix (Auction.h:273,283).
 * prices_out (nullable): cols doubles, the final object prices.  opts, stats nullable.
 * The fp32 entry computes a - p in double on the exactly widened coefficients, i.e. it is
 * Auction<double> on the widened matrix, the instantiation that terminates.
 */
int b200lsap_solve_f32(b200lsap_handle* handle, const float* a, int64_t rows, int64_t cols, int64_t ld,
                       int layout, const b200lsap_options* opts, int64_t* row_to_col, double* prices_out,
                       b200lsap_stats* stats);
int b200lsap_solve_f64(b200lsap_handle* handle, const double* a, int64_t rows, int64_t cols, int64_t ld,
                       int layout, const b200lsap_options* opts, int64_t* row_to_col, double* prices_out,
                       b200lsap_stats* stats);

/* Same solve on a matrix already in device memory.  ROW_MAJOR with ld % 4 == 0 and a
 * 16-byte aligned base is used in place; anything else is first re-laid out on the device.
 * row_to_col_dev: int32[rows] in device memory.  prices_dev nullable (double[cols], device). */
int b200lsap_solve_device_f32(b200lsap_handle* handle, const float* a_dev, int64_t rows, int64_t cols,
                              int64_t ld, int layout, const b200lsap_options* opts, int32_t* row_to_col_dev,
                              double* prices_dev, b200lsap_stats* stats);
int b200lsap_solve_device_f64(b200lsap_handle* handle, const double* a_dev, int64_t rows, int64_t cols,
                              int64_t ld, int layout, const b200lsap_options* opts, int32_t* row_to_col_dev,
                              double* prices_dev, b200lsap_stats* stats);

/*
 * Replaces Auction<Scalar, Eigen::SparseMatrix<Scalar, Eigen::RowMajor>>::solve (findForward on the
 * stored entries of a row, Auction.h:166-198, under the controller of Auction.h:44-109; the
 * driver's second call, LSAP.cpp:36-38,51).  The matrix is compressed row-major: values /
 * col_idx / row_ptr are Eigen's valuePtr() / innerIndexPtr() / outerIndexPtr() of a compressed
 * matrix (StorageIndex int), row_ptr[rows] == nnz, columns ascending inside a row.  Every
 * stored entry is an edge (the reference's sparse scan skips nothing); a missing entry is no
 * edge.  row_to_col, prices_out, opts and stats as in b200lsap_solve_f32; eps scaling (square
 * problems) starts at options.eps_start or 0.25.  A row without a stored entry is reported as
 * B200LSAP_ERR_INFEASIBLE; a problem without a perfect matching ends with B200LSAP_ERR_MAX_ROUNDS
 * (the reference does not return in either case).  The _device variants take the three arrays
 * in device memory, int32 row_to_col_dev and nullable prices_dev in device memory.
 */
int b200lsap_solve_csr_f32(b200lsap_handle* handle, const float* values, const int32_t* row_ptr,
                           const int32_t* col_idx, int64_t rows, int64_t cols, int64_t nnz,
                           const b200lsap_options* opts, int64_t* row_to_col, double* prices_out,
                           b200lsap_stats* stats);
int b200lsap_solve_csr_f64(b200lsap_handle* handle, const double* values, const int32_t* row_ptr,
                           const int32_t* col_idx, int64_t rows, int64_t cols, int64_t nnz,
                           const b200lsap_options* opts, int64_t* row_to_col, double* prices_out,
                           b200lsap_stats* stats);
int b200lsap_solve_csr_device_f32(b200lsap_handle* handle, const float* values_dev, const int32_t* row_ptr_dev,
                                  const int32_t* col_idx_dev, int64_t rows, int64_t cols, int64_t nnz,
                                  const b200lsap_options* opts, int32_t* row_to_col_dev, double* prices_dev,
                                  b200lsap_stats* stats);
int b200lsap_solve_csr_device_f64(b200lsap_handle* handle, const double* values_dev, const int32_t* row_ptr_dev,
                                  const int32_t* col_idx_dev, int64_t rows, int64_t cols, int64_t nnz,
                                  const b200lsap_options* opts, int32_t* row_to_col_dev, double* prices_dev,
                                  b200lsap_stats* stats);

/*
 * Batched mode for MHT-style data association: `batch` independent rows x cols problems,
 * problem b at a + b*rows*cols, each stored in `layout` with ld = its minor dimension.
 * rows <= 32, rows <= cols <= 64.  The problems are solved in the reference's own Gauss-Seidel order
 * (Auction.h:210-306 called once per problem, eps = eps_multiplier / cols), so every problem gets exactly
 * the assignment Auction<double>::solve returns for its coefficients widened to double.  A warp works on
 * sixteen problems at a time (two lanes each; cols % 4 == 0 and a 16-byte aligned matrix pointer) or on
 * eight (four lanes each; any cols).  Coefficients must be finite.
 * row_to_col: int32[batch*rows].  Host and device-resident variants.
 */
int b200lsap_solve_batch_f32(b200lsap_handle* handle, const float* a, int64_t batch, int32_t rows,
                             int32_t cols, int layout, const b200lsap_options* opts, int32_t* row_to_col,
                             b200lsap_stats* stats);
int b200lsap_solve_batch_device_f32(b200lsap_handle* handle, const float* a_dev, int64_t batch, int32_t rows,
                                    int32_t cols, int layout, const b200lsap_options* opts,
                                    int32_t* row_to_col_dev, b200lsap_stats* stats);

/* Synthetic inputs (include/b200lsap_gen.h) written straight into device memory:
 * `batch` consecutive problems starting at problem id problem0, each rows x cols with leading
 * dimension ld in `layout`, problem stride = ld * (layout ? rows : cols) elements. */
int b200lsap_generate_device_f32(b200lsap_handle* handle, float* a_dev, int64_t batch, int64_t rows,
                                 int64_t cols, int64_t ld, int layout, uint64_t seed, uint64_t problem0);
/* Column shard of one big problem: fills local columns [0, cols_local) with the values of
 * global columns [col0, col0+cols_local) of a rows x cols_global matrix. */
int b200lsap_generate_shard_device_f32(b200lsap_handle* handle, float* a_dev, int64_t rows,
                                       int64_t cols_local, int64_t ld, int64_t col0, int64_t cols_global,
                                       uint64_t seed, uint64_t problem);

/*
 * Column-sharded solve of ONE big fp32 problem over several GPUs — AuctionMT's decomposition
 * (column intervals per worker, AuctionMT.h:76-77 and AuctionCommon.h:302-320; exact top-2 merge in
 * interval order, AuctionMT.h:377-398) with a GPU per interval and one exchange per round.
 * Every rank holds columns [col0, col0+cols_local) of the matrix (row-major, 16-byte aligned, ld
 * and col0 multiples of 4, ld >= cols_local rounded up to 4 with the columns in between zero;
 * ranks in increasing col0 order) and a replica of the prices
 * and of the assignment.  Per round, on every rank:
 *     shard_bid      -> n records (b200lsap_shard_record_bytes() each) in rec_dev
 *     all-gather of the records, rank-major: gathered[r*n + b]      (the caller's: NCCL, NVLink)
 *     shard_resolve  -> merges, resolves, updates the replica; n_next bidders remain
 * until n_next == 0; then shard_next_phase (eps scaling, square problems) and, when it reports
 * no more phases, shard_finish.  STREAM CONTRACT: shard_bid enqueues its kernel on the handle's stream and returns
 * without synchronising, and shard_resolve reads `gathered` on that stream: the caller's exchange must run on the SAME
 * stream (b200lsap_set_stream with the stream of its collective, e.g. torch's current stream — what
 * auctionalgorithm_b200.sharded.CudaShardEngine does) or be ordered against it with events.  begin returns this rank's max |a|; the caller must set the
 * maximum over all ranks with shard_set_amax before the first bid (it scales the slack of the
 * single-precision filter).  auctionalgorithm_b200/sharded.py drives this with torch.distributed.
 */
int64_t b200lsap_shard_record_bytes(void);
int b200lsap_shard_begin_f32(b200lsap_handle* handle, const float* a_dev, int64_t rows, int64_t cols_local,
                             int64_t ld, int64_t col0, int64_t cols_global, const b200lsap_options* opts,
                             double* amax_local);
int b200lsap_shard_set_amax(b200lsap_handle* handle, double amax_global);
int b200lsap_shard_bid(b200lsap_handle* handle, void* rec_dev, int64_t* n_out);
int b200lsap_shard_resolve(b200lsap_handle* handle, const void* gathered_dev, int n_ranks, int64_t* n_next);
int b200lsap_shard_next_phase(b200lsap_handle* handle, int* more);
int b200lsap_shard_finish(b200lsap_handle* handle, int32_t* row_to_col_dev, double* prices_dev,
                          b200lsap_stats* stats);

/*
 * Fused form of the column-sharded solve: the whole solve is ONE persistent cooperative kernel per rank and the
 * per-round exchange happens inside it — every rank stores its 32-byte records straight into every other rank's
 * mailbox over NVLink (peer memory) and raises a flag there; merge and resolve run on every rank's replica.  The
 * host is not touched per round and the library needs no communicator: the caller wires the mailboxes ONCE.
 *     shard_comm_create   allocates this rank's mailbox for problems of up to rows_cap rows and n_ranks (<= 16)
 *                         ranks; ipc_handle_out (nullable) receives its CUDA IPC handle (64 bytes)
 *     shard_comm_connect  rank / n_ranks and, for every rank, EITHER its IPC handle (ipc_handles: n_ranks * 64
 *                         bytes, rank-major; ranks in other processes, e.g. exchanged with an all-gather) OR its
 *                         mailbox pointer (peer_ptrs[q] = b200lsap_shard_comm_ptr of rank q's handle; ranks in this
 *                         process, peer access is enabled here).  ranks_on_this_device > 1 (tests: several shards
 *                         on one GPU) divides the grid so that the ranks' kernels are resident together.
 *     solve_sharded_device_f32   collective: every rank calls it with its shard (same layout rules as
 *                         shard_begin_f32, ranks in increasing col0 order = rank order); returns this rank's
 *                         replica of the assignment and of the prices.  _begin / _end split it so that one host
 *                         thread can start every rank of a single-process run before it waits for any; such a
 *                         caller first calls shard_reserve on every handle (device buffers for the problem size:
 *                         growing them later would free memory, which waits for the kernels already running).
 * A rank that never arrives ends the others with B200LSAP_ERR_CUDA after a spin limit (10 s; the mailboxes must
 * then be wired again) instead of hanging the GPU.  Same rounds as the host-driven form above: identical prices.
 */
#define B200LSAP_IPC_HANDLE_BYTES 64
int b200lsap_shard_comm_create(b200lsap_handle* handle, int64_t rows_cap, int n_ranks, void* ipc_handle_out);
void* b200lsap_shard_comm_ptr(b200lsap_handle* handle);
int b200lsap_shard_comm_connect(b200lsap_handle* handle, int rank, int n_ranks, const void* ipc_handles,
                                void* const* peer_ptrs, int ranks_on_this_device);
int b200lsap_shard_reserve(b200lsap_handle* handle, int64_t rows, int64_t cols_global);
int b200lsap_solve_sharded_begin_f32(b200lsap_handle* handle, const float* a_dev, int64_t rows, int64_t cols_local,
                                     int64_t ld, int64_t col0, int64_t cols_global, const b200lsap_options* opts);
int b200lsap_solve_sharded_end(b200lsap_handle* handle, int32_t* row_to_col_dev, double* prices_dev,
                               b200lsap_stats* stats);
int b200lsap_solve_sharded_device_f32(b200lsap_handle* handle, const float* a_dev, int64_t rows, int64_t cols_local,
                                      int64_t ld, int64_t col0, int64_t cols_global, const b200lsap_options* opts,
                                      int32_t* row_to_col_dev, double* prices_dev, b200lsap_stats* stats);

/*
 * Several GPUs from ONE host thread.  b200lsap_multi owns a handle per device (devices[] may name a device more
 * than once: several ranks then share it), enables peer access between them and wires the mailboxes of the fused
 * sharded solve, so the column-sharded and the batch-split mode need neither NCCL nor torch.distributed:
 *     multi_solve_f32        Auction<Scalar>::solve / AuctionMT<Scalar>::solve (Auction.h:44, AuctionMT.h:50) on a
 *                            HOST matrix, column intervals over the devices (AuctionMT.h:76-77,
 *                            AuctionCommon.h:302-320), one in-kernel exchange per round; same result as
 *                            b200lsap_solve_f32 with wide rounds only (identical prices)
 *     multi_solve_batch_f32  b200lsap_solve_batch_f32 with the batch cut into contiguous slices, one per device,
 *                            every device running its own upload / kernel / download pipeline
 * Arguments as in the single-device entries.  Errors: the status is returned, the text is in multi_last_error.
 */
typedef struct b200lsap_multi b200lsap_multi;
int b200lsap_create_multi(b200lsap_multi** multi, const int* devices, int ndev);
int b200lsap_destroy_multi(b200lsap_multi* multi);
const char* b200lsap_multi_last_error(const b200lsap_multi* multi);
int b200lsap_multi_devices(const b200lsap_multi* multi);
int b200lsap_multi_solve_f32(b200lsap_multi* multi, const float* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                             const b200lsap_options* opts, int64_t* row_to_col, double* prices_out,
                             b200lsap_stats* stats);
int b200lsap_multi_solve_batch_f32(b200lsap_multi* multi, const float* a, int64_t batch, int32_t rows, int32_t cols,
                                   int layout, const b200lsap_options* opts, int32_t* row_to_col,
                                   b200lsap_stats* stats);

/*
 * Allocation-free batched solve.  This entry point has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same problems, same kernels and same result as
 * b200lsap_solve_batch_device_f32, but without a handle and without any allocation: it calls no cudaMalloc,
 * cudaFree or allocating library routine.  Its scratch — the launch's control block and, for column-major problems
 * (layout 0), the staging area of the re-layout pass — is ONE file-scope device array of
 * B200LSAP_NOALLOC_SCRATCH_BYTES (8 MiB + 256 B) that every GPU which loads the library holds until exit.
 *   - a_dev, row_to_col_dev (int32[batch*rows]) and status_dev (nullable int32: 0, or the b200lsap_status of a failed
 *     problem, whose rows read -1) are DEVICE pointers owned by the caller; device < 0 = the current device;
 *     cuda_stream = the cudaStream_t to enqueue on (NULL: the legacy default stream).
 *   - asynchronous: it returns once the work is enqueued (nothing is copied to the host, nothing synchronises), so it
 *     may be captured into a CUDA graph; the result is ready when the stream reaches that point.
 *   - limit: a column-major batch may hold at most B200LSAP_NOALLOC_MAX_ELEMS (2 097 152) coefficients
 *     (batch * rows * cols; 1365 problems of 32 x 48); beyond it the call is refused with B200LSAP_ERR_SHAPE.
 *     Row-major problems (layout 1) use only the control block and have no such limit.
 *   - every call on a device shares the array: calls that OVERLAP on the GPU (two streams, two threads, a graph
 *     replayed beside another call) race on it — serialise them.  The control block is cleared on the stream at the
 *     start of every call; the staging area keeps the last call's contents and is overwritten before it is read.
 *   - opts (nullable): eps_multiplier, max_rounds (bid watchdog per problem) and batch_bid_cap are honoured.
 * Argument errors are returned at once; their text is in b200lsap_noalloc_last_error().
 */
#define B200LSAP_NOALLOC_MAX_ELEMS 2097152
#define B200LSAP_NOALLOC_SCRATCH_BYTES (256 + 4 * B200LSAP_NOALLOC_MAX_ELEMS)
int b200lsap_solve_batch_noalloc_f32(int device, void* cuda_stream, const float* a_dev, int64_t batch, int32_t rows,
                                     int32_t cols, int layout, const b200lsap_options* opts, int32_t* row_to_col_dev,
                                     int32_t* status_dev);
const char* b200lsap_noalloc_last_error(void);

/* Full-sweep bidding pass for measurement: every row bids once, nothing is updated.
 * use_last_prices = 0: the state a solve starts from (zero prices, no candidate lists: the first
 * sweep of the first eps phase).  1: the state the handle's last dense solve left: its prices and,
 * when that solve ran in place on this very device matrix (same pointer, shape and ld; the
 * caller has not changed it), the rows' candidate lists, i.e. what the first sweep of every
 * later eps phase starts from.  Returns through ms_out the CUDA-event time of that single
 * kernel; bytes_out = rows*cols*sizeof(float). */
int b200lsap_bid_sweep_device_f32(b200lsap_handle* handle, const float* a_dev, int64_t rows, int64_t cols,
                                  int64_t ld, int use_last_prices, double* ms_out, int64_t* bytes_out);

/* Developer aid: cycle counters of the narrow rounds / tail of the last dense solve; all zero
 * unless the library was built with -DB200LSAP_TIMING.  out: up to 16 values. */
int b200lsap_debug_counters(const b200lsap_handle* handle, int64_t* out, int n);

/* Pinned host memory for callers that want full-rate PCIe copies. */
void* b200lsap_host_alloc(int64_t bytes);
void b200lsap_host_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* B200LSAP_H_ */
