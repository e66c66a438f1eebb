/*
 * TEST INFRASTRUCTURE — CPU restatement of the reference's dense auction path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * use anything under oracle/.  The product (auctionalgorithm_b200/) never links or calls it.
 *
 * This file is included twice by auction_oracle.c, once per arithmetic type:
 *   ORC_T      the Scalar (float or double) — every add, subtract and compare is done in it
 *   ORC_NAME   suffix of the generated functions
 *
 * What is restated (all line numbers relative to /root/reference/AuctionAlgorithm/):
 *   orc_best_two_*      Auction.h:124-152   findForward, dense: strict '>' top-2 over one row,
 *                                           a == 0 skipped, second best may equal best on ties
 *                       Auction.h:166-198   findForward, sparse row-major (CSR): the same over the
 *                                           stored entries of the row, in storage order, none skipped
 *   orc_forward_sweep_* Auction.h:210-306   forward: one Gauss-Seidel sweep over unlocked rows;
 *                                           bid = a - w + eps (:250), profit = w - eps (:253),
 *                                           lambda test (:257), edge-list eviction scan (:264-285)
 *   orc_best_two_col_*  Auction.h:308-334   findReverse (a.coeff(i, j): for CSR a search in row i)
 *   orc_reverse_sweep_* Auction.h:346-469   reverse, including the lambda rescaling of case 2
 *   orc_solve_*         Auction.h:44-109    controller: lambda = 0, eps = 1e-5 / cols
 *                                           (internal.h:16), prices 0, profits 1, steps 1-3
 *   termination tests   AuctionCommon.h:347-354 (unassignedObjectsLTlambda), :360-366
 *
 * Differences from the reference, none of which change a result:
 *   - plain C arrays instead of std::vector / vector<bool>;
 *   - a bid budget (max_bids) so a bounded sample can be timed and float livelocks
 *      end: when it runs out the solve stops with status ORC_BUDGET;
 *   - counters.
 */

#define ORC_CAT2(a, b) a##b
#define ORC_CAT(a, b) ORC_CAT2(a, b)
#define ORC_FN(base) ORC_CAT(base, ORC_NAME)

typedef struct
{
    const ORC_T* a;
    int64_t rows, cols;
    int64_t rs, cs; /* element strides: a(i,j) = a[i*rs + j*cs] */
    const int32_t* csr_ptr; /* non-NULL: a holds the stored values of a compressed row-major sparse */
    const int32_t* csr_col; /* matrix (Eigen outerIndexPtr / innerIndexPtr), columns ascending in a row */
    ORC_T* prices;
    ORC_T* profits;
    unsigned char* row_locked;
    unsigned char* col_locked;
    int64_t* ex;
    int64_t* ey;
    ORC_T* ev;
    int64_t n_edges;
    ORC_T lambda, epsilon;
    int64_t bids_left;
    oracle_stats* st;
} ORC_FN(orc_state_);

/* Auction.h:124-152 */
static int ORC_FN(orc_best_two_)(const ORC_FN(orc_state_) * s, int64_t i, ORC_T* v_i, ORC_T* w_i,
                                 ORC_T* a_best, int64_t* j_i)
{
    int found = 0;
    if (s->csr_ptr)
    {
        /* Auction.h:166-198 */
        for (int32_t k = s->csr_ptr[i]; k < s->csr_ptr[i + 1]; k++)
        {
            const int64_t j = s->csr_col[k];
            const ORC_T aij = s->a[k];
            const ORC_T diff = aij - s->prices[j];
            if (diff > *v_i)
            {
                if (found)
                    *w_i = *v_i;
                *v_i = diff;
                *j_i = j;
                *a_best = aij;
                found = 1;
            }
            if (diff > *w_i && *j_i != j)
                *w_i = diff;
        }
        return found;
    }
    const ORC_T* row = s->a + i * s->rs;
    const int64_t cs = s->cs;
    for (int64_t j = 0; j < s->cols; j++)
    {
        const ORC_T aij = row[j * cs];
        if (aij == 0)
            continue;
        const ORC_T diff = aij - s->prices[j];
        if (diff > *v_i)
        {
            if (found)
                *w_i = *v_i;
            *v_i = diff;
            *j_i = j;
            *a_best = aij;
            found = 1;
        }
        if (diff > *w_i && *j_i != j)
            *w_i = diff;
    }
    return found;
}

/* Auction.h:210-306; returns 1 if any assignment was made in this sweep */
static int ORC_FN(orc_forward_sweep_)(ORC_FN(orc_state_) * s)
{
    int any = 0;
    s->st->sweeps++;
    for (int64_t i = 0; i < s->rows; i++)
    {
        if (s->row_locked[i])
            continue;
        int64_t j_i = 0;
        ORC_T w_i = (ORC_T)-ORACLE_INF, v_i = (ORC_T)-ORACLE_INF, a_best = 0;
        if (!ORC_FN(orc_best_two_)(s, i, &v_i, &w_i, &a_best, &j_i))
            continue;
        if (s->bids_left == 0)
            return -1;
        if (s->bids_left > 0)
            s->bids_left--;
        s->st->bids++;

        const ORC_T bid = a_best - w_i + s->epsilon;
        s->profits[i] = w_i - s->epsilon;
        if (s->lambda <= bid)
        {
            s->prices[j_i] = bid;
            s->row_locked[i] = 1;
            s->col_locked[j_i] = 1;
            int fresh = 1;
            for (int64_t e = 0; e < s->n_edges; e++)
            {
                s->st->scan_steps++;
                if (s->ey[e] == j_i)
                {
                    s->row_locked[s->ex[e]] = 0;
                    fresh = 0;
                    s->ex[e] = i;
                    s->ev[e] = a_best;
                    s->st->evictions++;
                    break;
                }
            }
            if (fresh)
            {
                s->ex[s->n_edges] = i;
                s->ey[s->n_edges] = j_i;
                s->ev[s->n_edges] = a_best;
                s->n_edges++;
            }
            any = 1;
        }
        else
        {
            s->prices[j_i] = s->lambda;
            s->st->lambda_rejects++;
        }
    }
    return any;
}

/* a.coeff(i, j) of the compressed sparse matrix: 0 when (i, j) is not stored */
static ORC_T ORC_FN(orc_csr_coeff_)(const ORC_FN(orc_state_) * s, int64_t i, int64_t j)
{
    int32_t lo = s->csr_ptr[i], hi = s->csr_ptr[i + 1];
    while (lo < hi)
    {
        const int32_t mid = lo + (hi - lo) / 2;
        if (s->csr_col[mid] < j)
            lo = mid + 1;
        else
            hi = mid;
    }
    return (lo < s->csr_ptr[i + 1] && s->csr_col[lo] == j) ? s->a[lo] : (ORC_T)0;
}

/* Auction.h:308-334 */
static int ORC_FN(orc_best_two_col_)(const ORC_FN(orc_state_) * s, int64_t j, ORC_T* b_j, ORC_T* g_j,
                                     ORC_T* a_best, int64_t* i_j)
{
    int found = 0;
    const ORC_T* col = s->csr_ptr ? NULL : s->a + j * s->cs;
    for (int64_t i = 0; i < s->rows; i++)
    {
        const ORC_T aij = s->csr_ptr ? ORC_FN(orc_csr_coeff_)(s, i, j) : col[i * s->rs];
        if (aij == 0)
            continue;
        const ORC_T diff = aij - s->profits[i];
        if (diff > *b_j)
        {
            if (found)
                *g_j = *b_j;
            *b_j = diff;
            *i_j = i;
            *a_best = aij;
            found = 1;
        }
        if (diff > *g_j && *i_j != i)
            *g_j = diff;
    }
    return found;
}

/* Auction.h:346-469 */
static int ORC_FN(orc_reverse_sweep_)(ORC_FN(orc_state_) * s)
{
    int any = 0;
    s->st->reverse_calls++;
    for (int64_t j = 0; j < s->cols; j++)
    {
        if (s->col_locked[j])
            continue;
        if (!(s->prices[j] > s->lambda))
            continue;
        int64_t i_j = 0;
        ORC_T b_j = (ORC_T)-ORACLE_INF, g_j = (ORC_T)-ORACLE_INF, a_best = 0;
        if (!ORC_FN(orc_best_two_col_)(s, j, &b_j, &g_j, &a_best, &i_j))
            continue;
        s->st->reverse_work++;
        if (b_j >= (s->lambda + s->epsilon))
        {
            const ORC_T diff = g_j - s->epsilon;
            const ORC_T mx = s->lambda > diff ? s->lambda : diff;
            s->prices[j] = mx;
            s->profits[i_j] = a_best - mx;
            s->row_locked[i_j] = 1;
            s->col_locked[j] = 1;
            int fresh = 1;
            for (int64_t e = 0; e < s->n_edges; e++)
                if (s->ex[e] == i_j)
                {
                    s->col_locked[s->ey[e]] = 0;
                    fresh = 0;
                    s->ey[e] = j;
                    s->ev[e] = a_best;
                    break;
                }
            if (fresh)
            {
                s->ex[s->n_edges] = i_j;
                s->ey[s->n_edges] = j;
                s->ev[s->n_edges] = a_best;
                s->n_edges++;
            }
            any = 1;
        }
        else
        {
            s->prices[j] = b_j - s->epsilon;
            int64_t below = 0;
            ORC_T new_lambda = s->lambda;
            for (int64_t k = 0; k < s->cols; k++)
                if (s->prices[k] < s->lambda)
                {
                    below++;
                    if (s->prices[k] < new_lambda)
                        new_lambda = s->prices[k];
                }
            if (below >= (s->cols - s->rows))
                s->lambda = new_lambda;
        }
    }
    return any;
}

/* AuctionCommon.h:360-366 */
static int ORC_FN(orc_all_rows_locked_)(const ORC_FN(orc_state_) * s)
{
    for (int64_t i = 0; i < s->rows; i++)
        if (!s->row_locked[i])
            return 0;
    return 1;
}

/* AuctionCommon.h:347-354 */
static int ORC_FN(orc_free_cols_at_most_lambda_)(const ORC_FN(orc_state_) * s)
{
    for (int64_t j = 0; j < s->cols; j++)
        if (!s->col_locked[j] && s->prices[j] > s->lambda)
            return 0;
    return 1;
}

/* Auction.h:44-109 */
static int ORC_FN(orc_solve_)(const ORC_T* a, const int32_t* csr_ptr, const int32_t* csr_col, int64_t rows,
                              int64_t cols, int64_t ld, int layout, int64_t max_bids, int64_t* edge_x,
                              int64_t* edge_y, ORC_T* edge_v, int64_t* n_edges, ORC_T* prices_out,
                              ORC_T* profits_out, oracle_stats* stats)
{
    oracle_stats local;
    ORC_FN(orc_state_) s;
    memset(&local, 0, sizeof local);
    s.st = stats ? stats : &local;
    memset(s.st, 0, sizeof *s.st);
    *n_edges = 0;
    if (rows <= 0 || cols <= 0)
        return ORC_OK;
    s.a = a;
    s.csr_ptr = csr_ptr;
    s.csr_col = csr_col;
    s.rows = rows;
    s.cols = cols;
    s.rs = layout == ORACLE_ROW_MAJOR ? ld : 1;
    s.cs = layout == ORACLE_ROW_MAJOR ? 1 : ld;
    s.prices = (ORC_T*)calloc((size_t)cols, sizeof(ORC_T));
    s.profits = (ORC_T*)malloc((size_t)rows * sizeof(ORC_T));
    s.row_locked = (unsigned char*)calloc((size_t)rows, 1);
    s.col_locked = (unsigned char*)calloc((size_t)cols, 1);
    s.ex = edge_x;
    s.ey = edge_y;
    s.ev = edge_v;
    s.n_edges = 0;
    s.lambda = 0;
    s.epsilon = (ORC_T)(ORACLE_EPS_MULT / (double)cols); /* double division, then narrowed */
    s.bids_left = max_bids;
    for (int64_t i = 0; i < rows; i++)
        s.profits[i] = 1;

    int status = ORC_OK;
    for (;;)
    {
        int r;
        while ((r = ORC_FN(orc_forward_sweep_)(&s)) > 0)
            ;
        if (r < 0)
        {
            status = ORC_BUDGET;
            break;
        }
        if (!ORC_FN(orc_all_rows_locked_)(&s))
        {
            /* step 2; its loop count is bounded here because the reference spins forever
             * when a row can never be assigned */
            int64_t guard = 0;
            while (!ORC_FN(orc_reverse_sweep_)(&s) || !ORC_FN(orc_free_cols_at_most_lambda_)(&s))
                if (++guard > 4 * (rows + cols))
                {
                    status = ORC_INFEASIBLE;
                    break;
                }
            if (status != ORC_OK)
                break;
        }
        if (ORC_FN(orc_all_rows_locked_)(&s))
        {
            for (;;)
            {
                ORC_FN(orc_reverse_sweep_)(&s);
                if (ORC_FN(orc_free_cols_at_most_lambda_)(&s))
                    break;
            }
            break;
        }
    }
    *n_edges = s.n_edges;
    if (prices_out)
        memcpy(prices_out, s.prices, (size_t)cols * sizeof(ORC_T));
    if (profits_out)
        memcpy(profits_out, s.profits, (size_t)rows * sizeof(ORC_T));
    free(s.prices);
    free(s.profits);
    free(s.row_locked);
    free(s.col_locked);
    s.st->status = status;
    return status;
}

int ORC_FN(oracle_auction_)(const ORC_T* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                            int64_t max_bids, int64_t* edge_x, int64_t* edge_y, ORC_T* edge_v,
                            int64_t* n_edges, ORC_T* prices_out, ORC_T* profits_out, oracle_stats* stats)
{
    return ORC_FN(orc_solve_)(a, NULL, NULL, rows, cols, ld, layout, max_bids, edge_x, edge_y, edge_v, n_edges,
                              prices_out, profits_out, stats);
}

/* Auction<Scalar, Eigen::SparseMatrix<Scalar, Eigen::RowMajor>>::solve on a compressed matrix:
 * values / col_idx / row_ptr = valuePtr() / innerIndexPtr() / outerIndexPtr() */
int ORC_FN(oracle_auction_csr_)(const ORC_T* values, const int32_t* row_ptr, const int32_t* col_idx,
                                int64_t rows, int64_t cols, int64_t max_bids, int64_t* edge_x,
                                int64_t* edge_y, ORC_T* edge_v, int64_t* n_edges, ORC_T* prices_out,
                                ORC_T* profits_out, oracle_stats* stats)
{
    return ORC_FN(orc_solve_)(values, row_ptr, col_idx, rows, cols, 0, ORACLE_ROW_MAJOR, max_bids, edge_x, edge_y,
                              edge_v, n_edges, prices_out, profits_out, stats);
}

#undef ORC_CAT2
#undef ORC_CAT
#undef ORC_FN
