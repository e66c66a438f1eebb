// capi.cu — libb200lsap.so: handle, device buffers and the extern "C" entry points declared in
// include/b200lsap.h.  Host-side controller work that the reference does in
// Auction<Scalar>::solve (Auction.h:44-58: eps, initial prices, locks) is done here; the rounds
// themselves run in dense_kernels.cuh / batch_kernels.cuh.  There is no CPU solve path.
#include "../../include/b200lsap.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "batch_kernels.cuh"
#include "dense_kernels.cuh"
#include "shard_kernels.cuh"
#include "sparse_kernels.cuh"

using namespace b200lsap;

namespace
{

struct DevBuf
{
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap)
            return cudaSuccess;
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
        bytes = (bytes + 255) & ~size_t(255);
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess)
            cap = bytes;
        return e;
    }
    void release()
    {
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <typename U> U* as() const { return static_cast<U*>(p); }
};

} // namespace

constexpr int kBatchSlots = 16; // control blocks of batched launches in flight on one handle (chunked host entry)

struct b200lsap_handle
{
    int device = 0;
    int num_sms = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    DevBuf a, staging, csr_ptr, csr_col, csr_val, prices, pricesf, cand, cache, cache_hdr, narrow, profits, slots, owner, row2col, bid_col, bid_val, list0, list1, partial, arrive,
        ctrl, out64, batch_in, batch_out, batch_ctrl;
    DenseCtrl* h_ctrl = nullptr; // pinned mirror
    BatchCtrl* h_bctrl = nullptr; // pinned mirrors of the batched launches' control blocks (kBatchSlots of them)
    cudaStream_t up = nullptr, down = nullptr; // host entry of the batched mode: H2D / D2H beside the kernels
    cudaEvent_t ev_up[16] = {}, ev_done[16] = {};
    std::map<const void*, int> occupancy; // kernel -> resident CTAs per SM (attributes set once per handle)
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    int64_t last_cols = 0; // prices of the last dense solve are valid for this many columns
    const void* last_a = nullptr; // matrix of that solve when it was used in place (candidate lists stay valid)
    int64_t last_rows = 0, last_ld = 0;
    int min_seg = 512;
    int sweep_split = 0;
    int max_parts = 4;
    int64_t launches = 0;
    double eps_probe_scale = 0.125; // developer knob B200LSAP_EPS_SCALE
    // column-sharded solve in progress (b200lsap_shard_*)
    struct
    {
        bool active = false;
        const float* A = nullptr;
        int64_t rows = 0, cols_local = 0, ld = 0, col0 = 0, cols_global = 0;
        std::vector<double> eps;
        int phase = 0, cur = 0;
        int64_t n = 0, rounds = 0, bids = 0, max_rounds = 0;
        int grid = 0;
    } sh;
    // fused sharded solve (b200lsap_shard_comm_* / b200lsap_solve_sharded_*): this rank's mailbox and its peers'
    struct
    {
        DevBuf box;
        ShardComm cm{};
        bool connected = false;
        int share = 1;                 // ranks that run on this GPU (tests: several shards on one device)
        std::vector<void*> opened;     // CUDA IPC mappings of the peers' mailboxes
        bool running = false;
        int64_t rows = 0, cols_global = 0, cols_local = 0;
        double eps_final = 0;
        int phases = 0;
    } fused;
};

namespace
{

int fail(b200lsap_handle* h, int status, const char* what, cudaError_t e = cudaSuccess)
{
    if (h)
    {
        h->err = what;
        if (e != cudaSuccess)
        {
            h->err += ": ";
            h->err += cudaGetErrorString(e);
        }
    }
    return status;
}

#define CU(call)                                                                                                   \
    do                                                                                                             \
    {                                                                                                              \
        cudaError_t e_ = (call);                                                                                   \
        if (e_ != cudaSuccess)                                                                                     \
            return fail(h, B200LSAP_ERR_CUDA, #call, e_);                                                          \
    } while (0)

inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

inline int grid_for(const b200lsap_handle* h, int64_t work_items, int per_block)
{
    int64_t g = (work_items + per_block - 1) / per_block;
    const int64_t cap = (int64_t)h->num_sms * 8;
    if (g > cap)
        g = cap;
    if (g < 1)
        g = 1;
    return (int)g;
}

} // namespace

extern "C" {

const char* b200lsap_status_string(int status)
{
    switch (status)
    {
    case B200LSAP_OK: return "ok";
    case B200LSAP_ERR_SHAPE: return "bad shape: need 1 <= rows <= cols and ld >= minor dimension";
    case B200LSAP_ERR_INFEASIBLE: return "infeasible: a row has no non-zero coefficient";
    case B200LSAP_ERR_MAX_ROUNDS: return "round watchdog expired";
    case B200LSAP_ERR_CUDA: return "CUDA error";
    case B200LSAP_ERR_ARG: return "bad argument";
    case B200LSAP_ERR_NO_DEVICE: return "no CUDA device";
    default: return "unknown status";
    }
}

const char* b200lsap_last_error(const b200lsap_handle* h) { return h ? h->err.c_str() : "null handle"; }

void b200lsap_options_default(b200lsap_options* o)
{
    if (!o)
        return;
    memset(o, 0, sizeof *o);
    o->eps_multiplier = 1e-5;
    o->eps_scaling = -1;
    o->eps_start = 0.0;
    o->eps_factor = 0.0;
    o->max_rounds = 0;
    o->driver = 0;
}

int b200lsap_create(b200lsap_handle** out, int device)
{
    if (!out)
        return B200LSAP_ERR_ARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0)
        return B200LSAP_ERR_NO_DEVICE;
    if (device < 0 && cudaGetDevice(&device) != cudaSuccess)
        return B200LSAP_ERR_NO_DEVICE;
    if (device >= n)
        return B200LSAP_ERR_ARG;
    b200lsap_handle* h = new b200lsap_handle();
    h->device = device;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMallocHost((void**)&h->h_ctrl, sizeof(DenseCtrl)) != cudaSuccess ||
        cudaMallocHost((void**)&h->h_bctrl, kBatchSlots * sizeof(BatchCtrl)) != cudaSuccess ||
        cudaStreamCreateWithFlags(&h->up, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&h->down, cudaStreamNonBlocking) != cudaSuccess)
    {
        delete h;
        return B200LSAP_ERR_CUDA;
    }
    h->own_stream = true;
    for (auto& e : h->ev)
        cudaEventCreate(&e);
    for (int k = 0; k < kBatchSlots; k++)
    {
        cudaEventCreateWithFlags(&h->ev_up[k], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&h->ev_done[k], cudaEventDisableTiming);
    }
    if (const char* s = getenv("B200LSAP_MAX_PARTS"))
    {
        int v = atoi(s);
        if (v >= 1 && v <= 64)
            h->max_parts = v;
    }
    if (const char* s = getenv("B200LSAP_EPS_SCALE"))
    {
        const double v = atof(s);
        if (v > 0)
            h->eps_probe_scale = v;
    }
    if (const char* s = getenv("B200LSAP_SWEEP_SPLIT"))
        h->sweep_split = std::max(0, atoi(s));
    if (const char* s = getenv("B200LSAP_MIN_SEG"))
    {
        int v = atoi(s);
        if (v >= 128)
            h->min_seg = (v + 127) & ~127;
    }
    *out = h;
    return B200LSAP_OK;
}

int b200lsap_destroy(b200lsap_handle* h)
{
    if (!h)
        return B200LSAP_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    for (DevBuf* b : {&h->a, &h->staging, &h->csr_ptr, &h->csr_col, &h->csr_val, &h->prices, &h->pricesf, &h->cand, &h->cache, &h->cache_hdr, &h->narrow, &h->profits, &h->slots, &h->owner, &h->row2col, &h->bid_col,
                      &h->bid_val, &h->list0, &h->list1, &h->partial, &h->arrive, &h->ctrl, &h->out64, &h->batch_in,
                      &h->batch_out, &h->batch_ctrl})
        b->release();
    for (void* p : h->fused.opened)
        cudaIpcCloseMemHandle(p);
    h->fused.box.release();
    for (auto& e : h->ev)
        if (e)
            cudaEventDestroy(e);
    if (h->h_ctrl)
        cudaFreeHost(h->h_ctrl);
    if (h->h_bctrl)
        cudaFreeHost(h->h_bctrl);
    for (int k = 0; k < kBatchSlots; k++)
    {
        if (h->ev_up[k])
            cudaEventDestroy(h->ev_up[k]);
        if (h->ev_done[k])
            cudaEventDestroy(h->ev_done[k]);
    }
    if (h->up)
        cudaStreamDestroy(h->up);
    if (h->down)
        cudaStreamDestroy(h->down);
    if (h->own_stream && h->stream)
        cudaStreamDestroy(h->stream);
    delete h;
    return B200LSAP_OK;
}

int b200lsap_set_stream(b200lsap_handle* h, void* s)
{
    if (!h)
        return B200LSAP_ERR_ARG;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    if (h->own_stream && h->stream)
        cudaStreamDestroy(h->stream);
    h->stream = static_cast<cudaStream_t>(s);
    h->own_stream = false;
    return B200LSAP_OK;
}

/* Cycle counters of the last persistent solve (all zero unless built with -DB200LSAP_TIMING). */
int b200lsap_debug_counters(const b200lsap_handle* h, int64_t* out, int n)
{
    if (!h || !out)
        return B200LSAP_ERR_ARG;
    for (int k = 0; k < n && k < 16; k++)
        out[k] = h->h_ctrl->dbg[k];
    return B200LSAP_OK;
}

void* b200lsap_host_alloc(int64_t bytes)
{
    void* p = nullptr;
    if (bytes <= 0 || cudaMallocHost(&p, (size_t)bytes) != cudaSuccess)
        return nullptr;
    return p;
}

void b200lsap_host_free(void* p)
{
    if (p)
        cudaFreeHost(p);
}

} // extern "C"

// =============================================================================================
// Dense solve
// =============================================================================================
namespace
{

struct DensePlan
{
    int64_t rows, cols, ldp;
    std::vector<double> eps;
    int64_t max_rounds;
    int driver;
    int tail_rows;
    int narrow_rows;
    bool eps_auto;     // first eps of the scaling schedule from the data (k_probe / k_schedule; fp32 only)
    double eps_factor;
    int use_cache;
    int asym_phases = 0;               // rows < cols: the last this-many phases end with the price repair (run_dense)
    const double* prices_in = nullptr; // warm start (options.prices_in): host pointer on the host entries, device otherwise
    bool prices_on_host = false;
};

// Warm start: the largest imported price bounds every price the filter will see (ctrl->pmax); NaNs and negative
// prices are refused through the status.
__global__ void __launch_bounds__(256) k_price_max(const double* __restrict__ prices, long long cols, DenseCtrl* ctrl)
{
    double m = 0.0;
    bool bad = false;
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < cols; j += (long long)gridDim.x * blockDim.x)
    {
        const double p = prices[j];
        bad = bad || !(p >= 0.0) || !(p < 1e300);
        m = fmax(m, p);
    }
    if (bad)
        ctrl->status = 5; // B200LSAP_ERR_ARG
    if (m > 0.0)
        atomicMax(reinterpret_cast<unsigned long long*>(&ctrl->pmax), (unsigned long long)__double_as_longlong(m));
}

// Post-condition of an asymmetric solve (AuctionCommon.h:347-354 with lambda = the lowest price of an assigned
// column): no unassigned column may be priced above an assigned one.  It holds by construction from zero prices;
// after a warm start it has to be checked.  out[0] = max price of an unassigned column, out[1] = min of an assigned.
__global__ void __launch_bounds__(256) k_lambda_check(const double* __restrict__ prices, const int* __restrict__ owner,
                                                      long long cols, unsigned long long* out)
{
    double hi = 0.0, lo = 1e300;
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < cols; j += (long long)gridDim.x * blockDim.x)
    {
        const double p = prices[j];
        if (owner[j] < 0)
            hi = fmax(hi, p);
        else
            lo = fmin(lo, p);
    }
    atomicMax(&out[0], (unsigned long long)__double_as_longlong(hi)); // non-negative doubles order like their bits
    atomicMin(&out[1], (unsigned long long)__double_as_longlong(lo));
}

// Continuation of a finished solve: the unassigned rows become the bidders of the last eps phase again.
// pass 0 (one thread): counters; pass 1 (grid): compaction.
template <typename T> __global__ void __launch_bounds__(kThreads, 1) k_resume(DenseState<T> s, int pass)
{
    DenseCtrl* c = s.ctrl;
    if (pass == 0)
    {
        c->count[0] = c->count[1] = 0;
        c->done = 0;
        c->phase = c->n_phases - 1;
        c->phase_work = 0;
        c->phase_budget = 0;
        s.narrow->arrived = 0u; // the scanner hand-shake of the row-scanning path starts from scratch, as after k_init
        return;
    }
    if (blockIdx.x == 0 && threadIdx.x < kNarrowMax)
        s.narrow->slot[threadIdx.x].flag = 0;
    const int cur = c->cur, lane = threadIdx.x & 31;
    const long long n_threads = (long long)gridDim.x * blockDim.x;
    const long long up = ((long long)s.rows + 31) & ~31ll;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < up; i += n_threads)
    {
        const bool push = i < s.rows && s.row2col[i] < 0;
        const unsigned m = __ballot_sync(kFullMask, push);
        if (m)
        {
            int base = 0;
            if (lane == 0)
                base = atomicAdd(&c->count[cur], __popc(m));
            base = __shfl_sync(kFullMask, base, 0);
            if (push)
                s.list[cur][base + __popc(m & ((1u << lane) - 1))] = (int)i;
        }
    }
}

// The collective move of the cols - rows "dummy bidders" that make an asymmetric problem square (each values every
// object the same, so it only ever wants the cheapest one): every price below `floor` rises to `floor`, and a row
// whose object was raised is unassigned and bids again.  Prices only go up, so every bound the scans keep (candidate
// lists, row caches, bid slots) stays valid; the rows that keep their object keep eps-CS (the others' prices rose).
// This is what the reference's reverse auction is for (Auction.h:346-469, lambda raised at :436-464), done from the
// side that keeps prices monotone.
__global__ void __launch_bounds__(256) k_floor_raise(double* __restrict__ prices, float* __restrict__ pricesf,
                                                     int* __restrict__ owner, int* __restrict__ row2col, long long cols,
                                                     double floor)
{
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < cols; j += (long long)gridDim.x * blockDim.x)
        if (prices[j] < floor)
        {
            prices[j] = floor;
            pricesf[j] = (float)floor;
            const int o = owner[j];
            if (o >= 0)
            {
                row2col[o] = -1;
                owner[j] = -1;
            }
        }
}

// owner[] as the inverse of the returned assignment (the narrow-round kernels keep their own copy while they run).
__global__ void __launch_bounds__(256) k_owner_from_assignment(const int* __restrict__ row2col, long long rows, long long cols,
                                                               int* __restrict__ owner, int pass)
{
    const long long n = pass == 0 ? cols : rows;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x)
    {
        if (pass == 0)
            owner[t] = -1;
        else if (row2col[t] >= 0)
            owner[row2col[t]] = (int)t;
    }
}

// eps schedule (the controller constants of Auction.h:54-55, plus optional scaling).
int make_plan(b200lsap_handle* h, int64_t rows, int64_t cols, const b200lsap_options* opts, DensePlan& plan)
{
    b200lsap_options o;
    b200lsap_options_default(&o);
    if (opts)
        o = *opts;
    if (!(o.eps_multiplier > 0))
        return fail(h, B200LSAP_ERR_ARG, "eps_multiplier must be positive");
    const double eps_final = o.eps_multiplier / (double)cols; // Auction.h:55
    // (a warm start keeps the eps schedule: measured on 2048 x 2048 with 2 % of the coefficients moved by 0.01, one
    // phase at the final eps from the previous frame's prices takes 38 939 rounds, the data-sized schedule 6 427
    // against 6 582 cold: imported prices save the price DISCOVERY of the early phases, not the phases themselves)
    bool scaling = o.eps_scaling < 0 ? (rows == cols) : (o.eps_scaling != 0);
    plan.prices_in = o.prices_in;
    plan.asym_phases = o.eps_scaling == 2 ? 1 : kMaxPhases;
    if (scaling && rows != cols && o.driver != 0)
        return fail(h, B200LSAP_ERR_ARG, "eps scaling on rows < cols needs the persistent driver (driver = 0)");
    plan.eps.clear();
    plan.eps_auto = scaling && !(o.eps_start > 0);
    plan.eps_factor = o.eps_factor > 1 ? o.eps_factor : 7.0;
    plan.use_cache = o.row_cache >= 0;
    if (scaling)
    {
        double e = o.eps_start > 0 ? o.eps_start : 0.25;
        const double f = plan.eps_factor;
        while (e > eps_final * 1.0000001 && (int)plan.eps.size() < kMaxPhases - 1)
        {
            plan.eps.push_back(e);
            e /= f;
        }
    }
    plan.eps.push_back(eps_final);
    plan.rows = rows;
    plan.cols = cols;
    plan.max_rounds = o.max_rounds > 0 ? o.max_rounds : 20000000;
    plan.driver = o.driver;
    if (o.tail_rows > kNarrowMax || o.narrow_rows > kNarrowMax)
        return fail(h, B200LSAP_ERR_ARG, "tail_rows and narrow_rows must be at most 256");
    plan.tail_rows = o.tail_rows == 0 ? 1 : (o.tail_rows < 0 ? 0 : o.tail_rows);
    plan.narrow_rows = o.narrow_rows == 0 ? kNarrowMax : (o.narrow_rows < 0 ? 0 : o.narrow_rows);
    return B200LSAP_OK;
}

template <typename T> int alloc_state(b200lsap_handle* h, int64_t rows, int64_t cols, int64_t ldp, int total_warps)
{
    CU(h->prices.ensure((ldp + 512) * sizeof(double)));
    CU(h->pricesf.ensure((ldp + 512) * sizeof(float)));
    CU(h->cand.ensure((size_t)rows * kCand * sizeof(int2)));
    CU(h->cache.ensure((size_t)rows * kCacheK * sizeof(CacheEnt<T>)));
    CU(h->cache_hdr.ensure((size_t)rows * sizeof(RowCacheHdr)));
    CU(h->narrow.ensure(sizeof(NarrowCtrl)));
    CU(h->profits.ensure(rows * sizeof(double)));
    CU(h->slots.ensure(cols * sizeof(BidSlot)));
    CU(h->owner.ensure(cols * sizeof(int)));
    CU(h->row2col.ensure(rows * sizeof(int)));
    CU(h->bid_col.ensure(rows * sizeof(int)));
    CU(h->bid_val.ensure(rows * sizeof(double)));
    CU(h->list0.ensure(rows * sizeof(int)));
    CU(h->list1.ensure(rows * sizeof(int)));
    CU(h->partial.ensure((size_t)std::max<int64_t>(total_warps + 64, 4 * rows) * sizeof(Top2<T>)));
    CU(h->arrive.ensure(rows * sizeof(int)));
    CU(h->ctrl.ensure(sizeof(DenseCtrl)));
    return B200LSAP_OK;
}

template <typename T> DenseState<T> bind_state(b200lsap_handle* h, const T* A, int64_t rows, int64_t cols, int64_t ld)
{
    DenseState<T> s;
    s.A = A;
    s.ld = ld;
    s.rows = (int)rows;
    s.cols = (int)cols;
    s.cols4 = (int)round_up(cols, 4);
    s.prices = h->prices.as<double>();
    s.pricesf = h->pricesf.as<float>();
    s.cand = h->cand.as<int2>();
    s.cache = h->cache.as<CacheEnt<T>>();
    s.cache_hdr = h->cache_hdr.as<RowCacheHdr>();
    s.use_cache = 0;
    s.narrow = h->narrow.as<NarrowCtrl>();
    s.narrow_rows = 0;
    s.max_parts = h->max_parts;
    s.profits = h->profits.as<double>();
    s.slots = h->slots.as<BidSlot>();
    s.owner = h->owner.as<int>();
    s.row2col = h->row2col.as<int>();
    s.bid_col = h->bid_col.as<int>();
    s.bid_val = h->bid_val.as<double>();
    s.list[0] = h->list0.as<int>();
    s.list[1] = h->list1.as<int>();
    s.partial = h->partial.as<Top2<T>>();
    s.arrive = h->arrive.as<int>();
    s.ctrl = h->ctrl.as<DenseCtrl>();
    s.min_seg = h->min_seg;
    s.sweep_split = h->sweep_split;
    s.tail_rows = 0;
    s.dry = 0;
    s.narrow_after = 0;
    s.phase_cut = 0;
    s.asym_repair = 0;
    static const int every = getenv("B200LSAP_CACHE_EVERY_PHASE") ? atoi(getenv("B200LSAP_CACHE_EVERY_PHASE")) : 0;
    s.rebuild_caches_every_phase = every;
    return s;
}

template <typename T> int persistent_grid(b200lsap_handle* h, int& grid)
{
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_auction_persistent<T>, kThreads, 0));
    if (per_sm < 1)
        return fail(h, B200LSAP_ERR_CUDA, "persistent kernel does not fit on an SM");
    grid = h->num_sms; // one CTA per SM
    return B200LSAP_OK;
}

// Runs the auction on a padded row-major device matrix.  On return h->row2col holds the
// assignment and h->h_ctrl the controller block.
template <typename T>
// resume: the state of the solve that has just finished on this handle is kept (prices, owners, caches); the rows
// that the caller has unassigned since (k_floor_raise) bid again in the LAST eps phase until every row is assigned.
int run_dense_once(b200lsap_handle* h, const T* A, int64_t ld, const DensePlan& plan, b200lsap_stats* stats,
                   bool resume = false)
{
    int grid = 0;
    int rc = persistent_grid<T>(h, grid);
    if (rc)
        return rc;
    const int total_warps = grid * kWarpsPerBlock;
    rc = alloc_state<T>(h, plan.rows, plan.cols, ld, total_warps);
    if (rc)
        return rc;
    DenseState<T> s = bind_state<T>(h, A, plan.rows, plan.cols, ld);
    s.tail_rows = plan.tail_rows;
    s.narrow_rows = plan.narrow_rows;

    DenseCtrl* hc = h->h_ctrl;
    const bool eps_auto = plan.eps_auto && !resume;
    s.use_cache = (plan.driver == 0 && plan.use_cache) ? 1 : 0;
    {
        // the cache-building sweep reads the matrix once (~6.5 TB/s); a small wide round costs ~20 us
        const double sweep_us = (double)plan.rows * (double)ld * sizeof(T) / 6.5e6;
        s.narrow_after = sweep_us > 1000.0 ? (int)std::min(1e6, sweep_us / 20.0) : 0;
    }
    if (const char* e = getenv("B200LSAP_PHASE_CUT"))
        s.phase_cut = plan.driver == 0 ? atoi(e) : 0;
    // rows < cols with scaled eps or imported prices: the price post-condition is kept at the end of every phase
    static const int repair_env = getenv("B200LSAP_ASYM_PHASES") ? atoi(getenv("B200LSAP_ASYM_PHASES")) : 0; // developer knob
    s.asym_repair = (plan.driver == 0 && plan.rows < plan.cols && (plan.prices_in || plan.eps.size() > 1 || resume))
                        ? (repair_env > 0 ? repair_env : plan.asym_phases)
                        : 0;
    std::vector<double> eps = plan.eps;
    if (resume)
    {
        eps.assign(1, plan.eps.back());
        k_resume<T><<<1, 1, 0, h->stream>>>(s, 0);
        k_resume<T><<<grid, kThreads, 0, h->stream>>>(s, 1);
        h->launches += 2;
    }
    else
    {
    memset(hc, 0, sizeof *hc);
    hc->count[0] = (int)plan.rows;
    hc->n_phases = (int)plan.eps.size();
    hc->max_rounds = plan.max_rounds;
    for (size_t k = 0; k < plan.eps.size(); k++)
        hc->eps[k] = plan.eps[k];
    hc->eps_auto = eps_auto ? 1 : 0;
    hc->eps_factor = plan.eps_factor;
    hc->eps_final = plan.eps.back();
    hc->eps_probe_scale = h->eps_probe_scale;
    hc->eps_cap = 0.25;
    hc->phase_budget = eps_auto ? 4 * plan.rows + 1024 : 0;
    CU(cudaMemcpyAsync(s.ctrl, hc, sizeof *hc, cudaMemcpyHostToDevice, h->stream));

    CU(cudaEventRecord(h->ev[1], h->stream));
    if (plan.prices_in)
    {
        // warm start: the caller's prices instead of zeros (Auction.h:58); every row starts unassigned, so
        // eps-complementary slackness holds for the empty assignment whatever the prices are
        CU(cudaMemcpyAsync(s.prices, plan.prices_in, plan.cols * sizeof(double),
                           plan.prices_on_host ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, h->stream));
    }
    k_init<T><<<grid, kThreads, 0, h->stream>>>(s, plan.prices_in ? 1 : 0);
    if (plan.prices_in)
    {
        k_price_max<<<grid_for(h, plan.cols, 256), 256, 0, h->stream>>>(s.prices, plan.cols, s.ctrl); // after k_init: it resets the status
        h->launches++;
    }
    k_absmax<T><<<grid_for(h, plan.rows * (int64_t)(s.cols4 / 4), 256 * 8), 256, 0, h->stream>>>(s.A, s.ld, s.rows, s.cols4,
                                                                                         s.ctrl);
    h->launches += 2;
    {
        if (eps_auto)
        {
            // first eps of the schedule from the data (dense_cache.cuh)
            const int n_probe = (int)std::min<int64_t>(kProbeRows, plan.rows);
            k_probe<T><<<n_probe, kThreads, 0, h->stream>>>(s);
            k_schedule<<<1, 32, 0, h->stream>>>(s.ctrl, n_probe);
            h->launches += 2;
            if (plan.driver != 0)
            {
                CU(cudaMemcpyAsync(hc, s.ctrl, sizeof *hc, cudaMemcpyDeviceToHost, h->stream));
                CU(cudaStreamSynchronize(h->stream));
                eps.assign(hc->eps, hc->eps + hc->n_phases);
            }
        }
    }
    }
    if (plan.driver == 0)
    {
        void* args[] = {&s};
        if (s.use_cache)
        {
            // row caches: per eps phase one cooperative launch (wide rounds, then a fresh cache
            // for every row) and one k_narrow launch (the narrow rounds on one SM, narrow_kernel.cuh),
            // all enqueued now; launches that find nothing left to do return at once.
            {
                int n_phases = (int)eps.size();
                if (eps_auto)
                {
                    CU(cudaMemcpyAsync(hc, s.ctrl, sizeof *hc, cudaMemcpyDeviceToHost, h->stream));
                    CU(cudaStreamSynchronize(h->stream));
                    n_phases = hc->n_phases;
                }
                const size_t dyn = (size_t)round_up(plan.cols, 2) * sizeof(double) + (size_t)plan.cols * sizeof(int);
                cudaFuncAttributes fa;
                CU(cudaFuncGetAttributes(&fa, k_narrow<true, T>));
                int optin = 0;
                CU(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
                const bool smem_prices = dyn + fa.sharedSizeBytes + 1024 <= (size_t)optin;
                if (smem_prices)
                    CU(cudaFuncSetAttribute(k_narrow<true, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
                // one batch of launches covers the schedule; only an escalated schedule (escalate_schedule)
                // needs another one
                for (int batch_no = 0; batch_no < 64; batch_no++)
                {
                    // (a repaired phase hands over to k_narrow once more: a launch pair extra per phase covers that;
                    // launches that find the solve done return at once)
                    for (int p = 0; p < n_phases * (s.asym_repair ? 2 : 1); p++)
                    {
                        CU(cudaLaunchCooperativeKernel((const void*)k_auction_persistent<T>, dim3(grid), dim3(kThreads),
                                                       args, 0, h->stream));
                        if (smem_prices)
                            k_narrow<true, T><<<1, kNarrowThreads, dyn, h->stream>>>(s);
                        else
                            k_narrow<false, T><<<1, kNarrowThreads, 0, h->stream>>>(s);
                        h->launches += 2;
                    }
                    CU(cudaEventRecord(h->ev[2], h->stream));
                    CU(cudaMemcpyAsync(hc, s.ctrl, sizeof *hc, cudaMemcpyDeviceToHost, h->stream));
                    CU(cudaStreamSynchronize(h->stream));
                    if (hc->done || hc->status != 0)
                        break;
                    n_phases = hc->n_phases - (hc->phase < 0 ? 0 : hc->phase);
                    if (n_phases < 1)
                        n_phases = 1;
                }
                if (!hc->done && hc->status == 0)
                    hc->status = B200LSAP_ERR_MAX_ROUNDS;
            }
        }
        else
        {
            CU(cudaLaunchCooperativeKernel((const void*)k_auction_persistent<T>, dim3(grid), dim3(kThreads), args, 0,
                                           h->stream));
            h->launches++;
        }
        CU(cudaEventRecord(h->ev[2], h->stream));
        CU(cudaMemcpyAsync(hc, s.ctrl, sizeof *hc, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
    }
    else
    {
        // host-driven rounds: same device functions, the controller loop of Auction.h:60-100 here
        int cur = 0, phase = 0;
        long long rounds = 0, bids = 0;
        for (;;)
        {
            CU(cudaMemcpyAsync(hc, s.ctrl, 4 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
            CU(cudaStreamSynchronize(h->stream));
            const int n = hc->count[cur];
            if (hc->status != 0)
                break;
            if (n == 0)
            {
                if (phase + 1 >= (int)eps.size())
                    break;
                phase++;
                k_reset<T><<<grid, kThreads, 0, h->stream>>>(s, cur);
                h->launches++;
                continue;
            }
            if (rounds >= plan.max_rounds)
            {
                hc->status = B200LSAP_ERR_MAX_ROUNDS;
                break;
            }
            k_bid<T><<<grid, kThreads, 0, h->stream>>>(s, cur, n, eps[phase]);
            k_resolve<T><<<grid, kThreads, 0, h->stream>>>(s, cur, n);
            h->launches += 2;
            cur ^= 1;
            rounds++;
            bids += n;
        }
        CU(cudaEventRecord(h->ev[2], h->stream));
        CU(cudaStreamSynchronize(h->stream));
        hc->rounds = rounds;
        hc->bids = bids;
        hc->gs_bids = hc->gs_episodes = hc->narrow_rounds = 0;
        hc->cache_hits = hc->cache_misses = 0;
        hc->phase = phase;
        hc->eps[0] = eps[0];
    }
    CU(cudaGetLastError());
    h->last_cols = plan.cols;
    h->last_a = nullptr;
    if (stats)
    {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]);
        stats->ms_solve = ms;
        stats->rounds = hc->rounds;
        stats->bids = hc->bids;
        stats->phases = hc->phase + 1;
        stats->narrow_rounds = hc->narrow_rounds;
        stats->tail_bids = hc->gs_bids;
        stats->tail_runs = hc->gs_episodes + hc->escalations;
        stats->bid_bytes = hc->bids * plan.cols * (int64_t)sizeof(T);
        stats->eps_final = plan.eps.back();
        stats->eps_first = hc->eps[0];
        stats->cache_hits = hc->cache_hits;
        stats->cache_misses = hc->cache_misses;
        stats->warm_fallback = hc->repairs; // floor raises at the ends of phases (rows < cols)
    }
    if (hc->status != 0)
        return fail(h, hc->status, b200lsap_status_string(hc->status));
    return B200LSAP_OK;
}

// rows < cols with eps scaling or a warm start: a forward auction alone may end with an unassigned object priced above
// an assigned one, which breaks the post-condition of AuctionCommon.h:347-354 and with it optimality.  The repair is
// the reverse auction's job in the reference (Auction.h:346-469); here: raise every price below the highest
// unassigned price F to F (k_floor_raise: prices stay monotone), unassign the rows whose object was raised, let them
// bid again at the final eps, and repeat until the condition holds.  Every iteration raises the floor, so it ends.
template <typename T>
int run_dense(b200lsap_handle* h, const T* A, int64_t ld, const DensePlan& plan, b200lsap_stats* stats)
{
    int rc = run_dense_once<T>(h, A, ld, plan, stats);
    if (rc || plan.rows == plan.cols || (!plan.prices_in && plan.eps.size() == 1))
        return rc; // square, or a cold fixed-eps solve: unassigned objects were never priced
    int64_t repairs = 0;
    for (int it = 0; it < 1000; it++)
    {
        CU(h->out64.ensure(2 * sizeof(unsigned long long)));
        unsigned long long init[2] = {0ull, (unsigned long long)__builtin_bit_cast(long long, 1e300)};
        CU(cudaMemcpyAsync(h->out64.p, init, sizeof init, cudaMemcpyHostToDevice, h->stream));
        for (int pass = 0; pass < 2; pass++)
            k_owner_from_assignment<<<grid_for(h, plan.cols, 256), 256, 0, h->stream>>>(h->row2col.as<int>(), plan.rows,
                                                                                     plan.cols, h->owner.as<int>(), pass);
        k_lambda_check<<<grid_for(h, plan.cols, 256), 256, 0, h->stream>>>(h->prices.as<double>(), h->owner.as<int>(),
                                                                        plan.cols, h->out64.as<unsigned long long>());
        unsigned long long got[2];
        CU(cudaMemcpyAsync(got, h->out64.p, sizeof got, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        h->launches += 3;
        if (got[0] <= got[1])
            break; // max price of an unassigned object <= min price of an assigned one
        if (plan.driver != 0)
            return fail(h, B200LSAP_ERR_ARG, "eps scaling / warm start on rows < cols needs the persistent driver (driver = 0)");
        const double floor = __builtin_bit_cast(double, (long long)got[0]);
        k_floor_raise<<<grid_for(h, plan.cols, 256), 256, 0, h->stream>>>(h->prices.as<double>(), h->pricesf.as<float>(),
                                                                       h->owner.as<int>(), h->row2col.as<int>(), plan.cols,
                                                                       floor);
        h->launches++;
        repairs++;
        b200lsap_stats more;
        memset(&more, 0, sizeof more);
        rc = run_dense_once<T>(h, A, ld, plan, &more, true);
        if (stats)
        {
            stats->ms_solve = more.ms_solve; // ev[1] of the first run .. ev[2] of the last
            stats->rounds = more.rounds;     // the controller block keeps counting
            stats->bids = more.bids;
            stats->narrow_rounds = more.narrow_rounds;
            stats->tail_bids = more.tail_bids;
            stats->cache_hits = more.cache_hits;
            stats->cache_misses = more.cache_misses;
            stats->bid_bytes = more.bid_bytes;
        }
        if (rc)
            return rc;
    }
    if (stats)
        stats->warm_fallback = h->h_ctrl->repairs + repairs; // in-kernel (per phase) + this safety net
    return B200LSAP_OK;
}

int check_shape(b200lsap_handle* h, const void* a, int64_t rows, int64_t cols, int64_t ld, int layout)
{
    if (!h)
        return B200LSAP_ERR_ARG;
    if (!a)
        return fail(h, B200LSAP_ERR_ARG, "null matrix");
    if (layout != B200LSAP_COL_MAJOR && layout != B200LSAP_ROW_MAJOR)
        return fail(h, B200LSAP_ERR_ARG, "layout must be 0 (column-major) or 1 (row-major)");
    if (rows < 1 || cols < 1 || rows > cols || cols > 0x7fff0000ll || ld < (layout ? cols : rows))
        return fail(h, B200LSAP_ERR_SHAPE, b200lsap_status_string(B200LSAP_ERR_SHAPE));
    return B200LSAP_OK;
}

// Brings the caller's matrix into the padded row-major device layout.  Returns the device
// pointer to scan (the caller's own buffer when it already has that layout).
template <typename T>
int stage_matrix(b200lsap_handle* h, const T* a, bool on_host, int64_t rows, int64_t cols, int64_t ld, int layout,
                 const T** A_out, int64_t* ld_out)
{
    const int64_t ldp = round_up(cols, 32);
    if (!on_host && layout == B200LSAP_ROW_MAJOR && ld % 4 == 0 && cols % 4 == 0 && ((uintptr_t)a & 15) == 0)
    {
        *A_out = a;
        *ld_out = ld;
        return B200LSAP_OK;
    }
    CU(h->a.ensure((size_t)rows * ldp * sizeof(T)));
    T* A = h->a.as<T>();
    if (layout == B200LSAP_ROW_MAJOR)
    {
        if (on_host)
        {
            CU(cudaMemcpy2DAsync(A, ldp * sizeof(T), a, ld * sizeof(T), cols * sizeof(T), rows, cudaMemcpyHostToDevice,
                                 h->stream));
            if (ldp > cols)
                CU(cudaMemset2DAsync(A + cols, ldp * sizeof(T), 0, (ldp - cols) * sizeof(T), rows, h->stream));
        }
        else
        {
            k_copy_pad<T><<<grid_for(h, rows * ldp, 256 * 8), 256, 0, h->stream>>>(a, ld, A, ldp, (int)rows, (int)cols);
            h->launches++;
        }
    }
    else
    {
        const T* src = a;
        int64_t ld_src = ld;
        if (on_host)
        {
            CU(h->staging.ensure((size_t)rows * cols * sizeof(T)));
            CU(cudaMemcpy2DAsync(h->staging.p, rows * sizeof(T), a, ld * sizeof(T), rows * sizeof(T), cols,
                                 cudaMemcpyHostToDevice, h->stream));
            src = h->staging.as<T>();
            ld_src = rows;
        }
        const int64_t tiles = ((rows + 31) / 32) * (ldp / 32);
        k_transpose_pad<T><<<grid_for(h, tiles, 1), 256, 0, h->stream>>>(src, ld_src, A, ldp, (int)rows, (int)cols);
        h->launches++;
    }
    CU(cudaGetLastError());
    *A_out = A;
    *ld_out = ldp;
    return B200LSAP_OK;
}

template <typename T>
int solve_host(b200lsap_handle* h, const T* a, int64_t rows, int64_t cols, int64_t ld, int layout,
               const b200lsap_options* opts, int64_t* row_to_col, double* prices_out, b200lsap_stats* stats)
{
    int rc = check_shape(h, a, rows, cols, ld, layout);
    if (rc)
        return rc;
    if (!row_to_col)
        return fail(h, B200LSAP_ERR_ARG, "null row_to_col");
    CU(cudaSetDevice(h->device));
    DensePlan plan;
    rc = make_plan(h, rows, cols, opts, plan);
    if (rc)
        return rc;
    if (stats)
        memset(stats, 0, sizeof *stats);
    h->launches = 0;
    CU(cudaEventRecord(h->ev[0], h->stream));
    const T* A = nullptr;
    int64_t ldA = 0;
    rc = stage_matrix<T>(h, a, true, rows, cols, ld, layout, &A, &ldA);
    if (rc)
        return rc;
    plan.prices_on_host = true;
    rc = run_dense<T>(h, A, ldA, plan, stats);
    if (rc)
        return rc;
    CU(h->out64.ensure(rows * sizeof(long long)));
    k_widen_i32<<<grid_for(h, rows, 256), 256, 0, h->stream>>>(h->row2col.as<int>(), h->out64.as<long long>(), rows);
    h->launches++;
    CU(cudaMemcpyAsync(row_to_col, h->out64.p, rows * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
    if (prices_out)
        CU(cudaMemcpyAsync(prices_out, h->prices.p, cols * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaEventRecord(h->ev[3], h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (stats)
    {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]);
        stats->ms_upload = ms;
        cudaEventElapsedTime(&ms, h->ev[2], h->ev[3]);
        stats->ms_download = ms;
        stats->launches = h->launches;
    }
    return B200LSAP_OK;
}

template <typename T>
int solve_device(b200lsap_handle* h, const T* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                 const b200lsap_options* opts, int32_t* row_to_col_dev, double* prices_dev, b200lsap_stats* stats)
{
    int rc = check_shape(h, a, rows, cols, ld, layout);
    if (rc)
        return rc;
    if (!row_to_col_dev)
        return fail(h, B200LSAP_ERR_ARG, "null row_to_col");
    CU(cudaSetDevice(h->device));
    DensePlan plan;
    rc = make_plan(h, rows, cols, opts, plan);
    if (rc)
        return rc;
    if (stats)
        memset(stats, 0, sizeof *stats);
    h->launches = 0;
    CU(cudaEventRecord(h->ev[0], h->stream));
    const T* A = nullptr;
    int64_t ldA = 0;
    rc = stage_matrix<T>(h, a, false, rows, cols, ld, layout, &A, &ldA);
    if (rc)
        return rc;
    rc = run_dense<T>(h, A, ldA, plan, stats);
    if (rc)
        return rc;
    if ((const void*)A == (const void*)a) // used in place: the candidate lists describe the caller's matrix
        h->last_a = a, h->last_rows = rows, h->last_ld = ld;
    CU(cudaMemcpyAsync(row_to_col_dev, h->row2col.p, rows * sizeof(int32_t), cudaMemcpyDeviceToDevice, h->stream));
    if (prices_dev)
        CU(cudaMemcpyAsync(prices_dev, h->prices.p, cols * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaEventRecord(h->ev[3], h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (stats)
    {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]);
        stats->ms_upload = ms;
        cudaEventElapsedTime(&ms, h->ev[2], h->ev[3]);
        stats->ms_download = ms;
        stats->launches = h->launches;
    }
    return B200LSAP_OK;
}

} // namespace

extern "C" {

int b200lsap_solve_f32(b200lsap_handle* h, const float* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                       const b200lsap_options* opts, int64_t* row_to_col, double* prices_out, b200lsap_stats* stats)
{
    return solve_host<float>(h, a, rows, cols, ld, layout, opts, row_to_col, prices_out, stats);
}

int b200lsap_solve_f64(b200lsap_handle* h, const double* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                       const b200lsap_options* opts, int64_t* row_to_col, double* prices_out, b200lsap_stats* stats)
{
    return solve_host<double>(h, a, rows, cols, ld, layout, opts, row_to_col, prices_out, stats);
}

int b200lsap_solve_device_f32(b200lsap_handle* h, const float* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                              const b200lsap_options* opts, int32_t* r2c, double* prices, b200lsap_stats* stats)
{
    return solve_device<float>(h, a, rows, cols, ld, layout, opts, r2c, prices, stats);
}

int b200lsap_solve_device_f64(b200lsap_handle* h, const double* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                              const b200lsap_options* opts, int32_t* r2c, double* prices, b200lsap_stats* stats)
{
    return solve_device<double>(h, a, rows, cols, ld, layout, opts, r2c, prices, stats);
}

} // extern "C"

// =============================================================================================
// Sparse row-major (CSR) solve: Auction<Scalar, Eigen::SparseMatrix<Scalar, Eigen::RowMajor>>::solve
// =============================================================================================
namespace
{

// start of a CSR solve: the part of k_init the sparse kernels need (no candidate lists, no row caches)
template <typename T> __global__ void __launch_bounds__(kThreads, 1) k_init_csr(DenseState<T> s)
{
    const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
    const long long n_threads = (long long)gridDim.x * kThreads;
    reset_phase(s, s.list[0], tid, n_threads);
    for (long long i = tid; i < s.rows; i += n_threads)
        s.profits[i] = 1.0; // Auction.h:58
    for (long long j = tid; j < s.cols; j += n_threads)
    {
        s.prices[j] = 0.0;
        s.pricesf[j] = 0.f;
        s.slots[j] = BidSlot{0ull, 0ull};
    }
}

// Validation of device-resident CSR arrays (one thread per row): status 5 (bad arrays) wins over 2 (empty row).
__global__ void __launch_bounds__(256) k_check_csr(const int* __restrict__ ptr, const int* __restrict__ col, long long rows,
                                                   long long cols, long long nnz, int* status)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < rows; i += (long long)gridDim.x * blockDim.x)
    {
        const long long a = ptr[i], b = ptr[i + 1];
        int bad = 0;
        if ((i == 0 && a != 0) || (i == rows - 1 && b != nnz) || b < a || a < 0 || b > nnz)
            bad = 5;
        else if (b == a)
            bad = 2;
        else
            for (long long k = a; k < b; k++)
                if (col[k] < 0 || col[k] >= cols || (k > a && col[k] <= col[k - 1]))
                    bad = 5;
        if (bad)
            atomicMax(status, bad);
    }
}

template <typename T>
int solve_csr(b200lsap_handle* h, const T* values, const int32_t* row_ptr, const int32_t* col_idx, int64_t rows,
              int64_t cols, int64_t nnz, bool on_host, const b200lsap_options* opts, void* row_to_col, double* prices_out,
              b200lsap_stats* stats)
{
    if (!h)
        return B200LSAP_ERR_ARG;
    if (!values || !row_ptr || !col_idx || !row_to_col)
        return fail(h, B200LSAP_ERR_ARG, "null argument");
    if (rows < 1 || cols < 1 || rows > cols || cols > 0x7fff0000ll || nnz < rows || nnz > 0x7fff0000ll)
        return fail(h, B200LSAP_ERR_SHAPE, b200lsap_status_string(B200LSAP_ERR_SHAPE));
    if (on_host)
    {
        // the reference never returns on these
        if (row_ptr[0] != 0 || row_ptr[rows] != nnz)
            return fail(h, B200LSAP_ERR_ARG, "row_ptr[0] must be 0 and row_ptr[rows] must equal nnz");
        for (int64_t i = 0; i < rows; i++)
        {
            if (row_ptr[i + 1] <= row_ptr[i])
                return fail(h, row_ptr[i + 1] < row_ptr[i] ? B200LSAP_ERR_ARG : B200LSAP_ERR_INFEASIBLE,
                            row_ptr[i + 1] < row_ptr[i] ? "row_ptr is not monotone" : "a row without a stored entry");
            // a column outside [0, cols) would index prices / owners / bid slots out of bounds on the device
            for (int64_t k = row_ptr[i]; k < row_ptr[i + 1]; k++)
                if (col_idx[k] < 0 || col_idx[k] >= cols || (k > row_ptr[i] && col_idx[k] <= col_idx[k - 1]))
                    return fail(h, B200LSAP_ERR_ARG,
                                "col_idx must lie in [0, cols) and ascend strictly inside a row (a compressed Eigen matrix)");
        }
    }
    CU(cudaSetDevice(h->device));
    DensePlan plan;
    int rc = make_plan(h, rows, cols, opts, plan);
    if (rc)
        return rc;
    if (stats)
        memset(stats, 0, sizeof *stats);
    h->launches = 0;
    CU(cudaEventRecord(h->ev[0], h->stream));
    CsrView<T> m;
    m.nnz = nnz;
    if (on_host)
    {
        CU(h->csr_ptr.ensure((rows + 1) * sizeof(int32_t)));
        CU(h->csr_col.ensure(nnz * sizeof(int32_t)));
        CU(h->csr_val.ensure(nnz * sizeof(T)));
        CU(cudaMemcpyAsync(h->csr_ptr.p, row_ptr, (rows + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
        CU(cudaMemcpyAsync(h->csr_col.p, col_idx, nnz * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
        CU(cudaMemcpyAsync(h->csr_val.p, values, nnz * sizeof(T), cudaMemcpyHostToDevice, h->stream));
        m.ptr = h->csr_ptr.as<int>();
        m.col = h->csr_col.as<int>();
        m.val = h->csr_val.as<T>();
    }
    else
    {
        m.ptr = row_ptr;
        m.col = col_idx;
        m.val = values;
        // device arrays: the same checks as on the host path, before anything is indexed with them
        CU(h->ctrl.ensure(sizeof(DenseCtrl)));
        int* flag = &h->ctrl.as<DenseCtrl>()->status;
        CU(cudaMemsetAsync(flag, 0, sizeof(int), h->stream));
        k_check_csr<<<grid_for(h, rows, 256), 256, 0, h->stream>>>(row_ptr, col_idx, rows, cols, nnz, flag);
        CU(cudaMemcpyAsync(&h->h_ctrl->status, flag, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        CU(cudaGetLastError());
        h->launches++;
        if (h->h_ctrl->status != 0)
            return fail(h, h->h_ctrl->status,
                        h->h_ctrl->status == B200LSAP_ERR_INFEASIBLE
                            ? "a row without a stored entry"
                            : "bad CSR arrays: row_ptr[0] == 0, row_ptr[rows] == nnz, row_ptr monotone, col_idx in [0, cols) "
                              "ascending strictly inside a row");
    }
    // the arrays of a round (prices, owners, bid slots, bidder lists); no candidate lists or row caches
    CU(h->prices.ensure((cols + 512) * sizeof(double)));
    CU(h->pricesf.ensure((cols + 512) * sizeof(float)));
    CU(h->profits.ensure(rows * sizeof(double)));
    CU(h->slots.ensure(cols * sizeof(BidSlot)));
    CU(h->owner.ensure(cols * sizeof(int)));
    CU(h->row2col.ensure(rows * sizeof(int)));
    CU(h->bid_col.ensure(rows * sizeof(int)));
    CU(h->bid_val.ensure(rows * sizeof(double)));
    CU(h->list0.ensure(rows * sizeof(int)));
    CU(h->list1.ensure(rows * sizeof(int)));
    CU(h->ctrl.ensure(sizeof(DenseCtrl)));
    DenseState<T> s = bind_state<T>(h, (const T*)nullptr, rows, cols, cols);
    DenseCtrl* hc = h->h_ctrl;
    memset(hc, 0, sizeof *hc);
    hc->count[0] = (int)rows;
    hc->n_phases = (int)plan.eps.size();
    hc->max_rounds = plan.max_rounds;
    for (size_t k = 0; k < plan.eps.size(); k++)
        hc->eps[k] = plan.eps[k];
    CU(cudaMemcpyAsync(s.ctrl, hc, sizeof *hc, cudaMemcpyHostToDevice, h->stream));
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_auction_csr<T>, kThreads, 0));
    if (per_sm < 1)
        return fail(h, B200LSAP_ERR_CUDA, "sparse kernel does not fit on an SM");
    // no more CTAs than the first round has warps of work for
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(h->num_sms, (rows + kWarpsPerBlock - 1) / kWarpsPerBlock));
    CU(cudaEventRecord(h->ev[1], h->stream));
    k_init_csr<T><<<grid, kThreads, 0, h->stream>>>(s);
    k_absmax_csr<T><<<grid_for(h, nnz, 256 * 8), 256, 0, h->stream>>>(m.val, nnz, s.ctrl);
    void* args[] = {&s, &m};
    CU(cudaLaunchCooperativeKernel((void*)k_auction_csr<T>, dim3(grid), dim3(kThreads), args, 0, h->stream));
    h->launches += 3;
    CU(cudaMemcpyAsync(hc, s.ctrl, sizeof *hc, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaEventRecord(h->ev[2], h->stream));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaGetLastError());
    h->last_cols = 0;
    h->last_a = nullptr;
    if (stats)
    {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]);
        stats->ms_solve = ms;
        stats->rounds = hc->rounds;
        stats->bids = hc->bids;
        stats->phases = hc->phase + 1;
        stats->narrow_rounds = hc->narrow_rounds;
        stats->bid_bytes = hc->bids * (nnz / rows) * (int64_t)(sizeof(T) + sizeof(int32_t)); // average row
        stats->eps_final = plan.eps.back();
        stats->eps_first = plan.eps.front();
    }
    if (hc->status != 0)
        return fail(h, hc->status, b200lsap_status_string(hc->status));
    if (on_host)
    {
        CU(h->out64.ensure(rows * sizeof(long long)));
        k_widen_i32<<<grid_for(h, rows, 256), 256, 0, h->stream>>>(h->row2col.as<int>(), h->out64.as<long long>(), rows);
        h->launches++;
        CU(cudaMemcpyAsync(row_to_col, h->out64.p, rows * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
        if (prices_out)
            CU(cudaMemcpyAsync(prices_out, h->prices.p, cols * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    }
    else
    {
        CU(cudaMemcpyAsync(row_to_col, h->row2col.p, rows * sizeof(int32_t), cudaMemcpyDeviceToDevice, h->stream));
        if (prices_out)
            CU(cudaMemcpyAsync(prices_out, h->prices.p, cols * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    }
    CU(cudaEventRecord(h->ev[3], h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (stats)
    {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]);
        stats->ms_upload = ms;
        cudaEventElapsedTime(&ms, h->ev[2], h->ev[3]);
        stats->ms_download = ms;
        stats->launches = h->launches;
    }
    return B200LSAP_OK;
}

} // namespace

extern "C" {

int b200lsap_solve_csr_f32(b200lsap_handle* h, const float* values, const int32_t* row_ptr, const int32_t* col_idx,
                           int64_t rows, int64_t cols, int64_t nnz, const b200lsap_options* opts, int64_t* row_to_col,
                           double* prices_out, b200lsap_stats* stats)
{
    return solve_csr<float>(h, values, row_ptr, col_idx, rows, cols, nnz, true, opts, row_to_col, prices_out, stats);
}

int b200lsap_solve_csr_f64(b200lsap_handle* h, const double* values, const int32_t* row_ptr, const int32_t* col_idx,
                           int64_t rows, int64_t cols, int64_t nnz, const b200lsap_options* opts, int64_t* row_to_col,
                           double* prices_out, b200lsap_stats* stats)
{
    return solve_csr<double>(h, values, row_ptr, col_idx, rows, cols, nnz, true, opts, row_to_col, prices_out, stats);
}

int b200lsap_solve_csr_device_f32(b200lsap_handle* h, const float* values, const int32_t* row_ptr, const int32_t* col_idx,
                                  int64_t rows, int64_t cols, int64_t nnz, const b200lsap_options* opts,
                                  int32_t* row_to_col_dev, double* prices_dev, b200lsap_stats* stats)
{
    return solve_csr<float>(h, values, row_ptr, col_idx, rows, cols, nnz, false, opts, row_to_col_dev, prices_dev, stats);
}

int b200lsap_solve_csr_device_f64(b200lsap_handle* h, const double* values, const int32_t* row_ptr, const int32_t* col_idx,
                                  int64_t rows, int64_t cols, int64_t nnz, const b200lsap_options* opts,
                                  int32_t* row_to_col_dev, double* prices_dev, b200lsap_stats* stats)
{
    return solve_csr<double>(h, values, row_ptr, col_idx, rows, cols, nnz, false, opts, row_to_col_dev, prices_dev, stats);
}

} // extern "C"

// =============================================================================================
// Batched small problems, generators, measurement sweep
// =============================================================================================
namespace
{

// Resident CTAs per SM of a batched kernel; its shared-memory attribute is set the first time a handle sees it
// (not on every call: a 100-problem MHT frame would pay two driver queries per solve).
int batch_occupancy(b200lsap_handle* h, const void* kernel, int threads, size_t smem, int& per_sm)
{
    auto it = h->occupancy.find(kernel);
    if (it != h->occupancy.end())
    {
        per_sm = it->second;
        return B200LSAP_OK;
    }
    if (smem > 0)
        CU(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
    h->occupancy[kernel] = per_sm;
    return B200LSAP_OK;
}

// After the stream has been synchronised: statistics and status of the launches that used slots [slot0, slot0 + n).
int batch_collect(b200lsap_handle* h, int slot0, int n, b200lsap_stats* stats)
{
    int status = 0;
    for (int k = slot0; k < slot0 + n; k++)
    {
        const BatchCtrl& c = h->h_bctrl[k];
        if (stats)
        {
            stats->batch_takeovers += c.takeovers;
            stats->batch_max_bids = std::max<int64_t>(stats->batch_max_bids, c.max_bids);
            stats->cache_hits += (int64_t)c.hits;     // takeover bids answered from a row cache
            stats->cache_misses += (int64_t)c.misses; // ... by a full scan of the row
        }
        for (int d = 0; d < 4; d++)
            h->h_ctrl->dbg[d] = (long long)c.dbg[d]; // b200lsap_debug_counters
        if (c.status != 0 && status == 0)
            status = c.status;
    }
    if (status != 0)
        return fail(h, status, b200lsap_status_string(status));
    return B200LSAP_OK;
}

// slot: which of the handle's control blocks the launch uses (launches of one call that are in flight together take
// different slots); sync = false: enqueue only, the caller synchronises and calls batch_collect.
int run_batch(b200lsap_handle* h, const float* a_dev_row_major, int64_t batch, int rows, int cols,
              const b200lsap_options* opts, int32_t* r2c_dev, b200lsap_stats* stats, bool sync = true, int slot = 0)
{
    b200lsap_options o;
    b200lsap_options_default(&o);
    if (opts)
        o = *opts;
    if (!(o.eps_multiplier > 0))
        return fail(h, B200LSAP_ERR_ARG, "eps_multiplier must be positive");
    BatchParams P;
    P.a = a_dev_row_major;
    P.row2col = r2c_dev;
    P.batch = batch;
    P.rows = rows;
    P.cols = cols;
    P.eps = o.eps_multiplier / (double)cols;
    P.max_bids = o.max_rounds > 0 ? (int)(o.max_rounds > 0x7fffffff ? 0x7fffffff : o.max_rounds) : 1000000;
    CU(h->batch_ctrl.ensure(kBatchSlots * sizeof(BatchCtrl)));
    P.ctl = h->batch_ctrl.as<BatchCtrl>() + slot;
    P.status = &P.ctl->status;
    // B200LSAP_BATCH_CAP / B200LSAP_BATCH_CHUNK: developer A/B knobs of the filtered kernel
    static const int cap_env = getenv("B200LSAP_BATCH_CAP") ? atoi(getenv("B200LSAP_BATCH_CAP")) : 0;
    static const int chunk_env = getenv("B200LSAP_BATCH_CHUNK") ? atoi(getenv("B200LSAP_BATCH_CHUNK")) : 0;
    P.cap = o.batch_bid_cap > 0 ? o.batch_bid_cap : (o.batch_bid_cap < 0 ? 0x7fffffff : (cap_env > 0 ? cap_env : 96));
    P.chunk = chunk_env > 0 ? chunk_env : 16;
    CU(cudaMemsetAsync(P.ctl, 0, sizeof(BatchCtrl), h->stream));
    using BatchKernel = void (*)(BatchParams);
    // B200LSAP_BATCH_SMEM=1: the shared-memory kernel (k_auction_batch); B200LSAP_BATCH_WARPS=16|24|32: warps per
    // SM the streaming kernel is compiled for (developer A/B knobs)
    static const bool smem_mode = getenv("B200LSAP_BATCH_SMEM") && atoi(getenv("B200LSAP_BATCH_SMEM")) != 0;
    static const int warps_mode = getenv("B200LSAP_BATCH_WARPS") ? atoi(getenv("B200LSAP_BATCH_WARPS")) : 32;
    // B200LSAP_BATCH_FILTER=0: without the single-precision filter (k_auction_batch_stream)
    static const bool filter_mode = !(getenv("B200LSAP_BATCH_FILTER") && atoi(getenv("B200LSAP_BATCH_FILTER")) == 0);
    CU(cudaEventRecord(h->ev[1], h->stream));
    if (!smem_mode && filter_mode && cols % 4 == 0 && ((uintptr_t)a_dev_row_major & 15) == 0)
    {
        // 16 problems per warp, four warps per CTA, one resident wave; warps reserve problems from one counter
        constexpr int kFW = 4, kFC = 4;
#define B200LSAP_FILTER(M) k_auction_batch_filter<M, true, kFW, kFC>, k_auction_batch_filter<M, false, kFW, kFC>
        static const BatchKernel kf[16] = {B200LSAP_FILTER(1), B200LSAP_FILTER(2), B200LSAP_FILTER(3),
                                           B200LSAP_FILTER(4), B200LSAP_FILTER(5), B200LSAP_FILTER(6),
                                           B200LSAP_FILTER(7), B200LSAP_FILTER(8)};
        const int kM = (cols + 7) / 8;
        const BatchKernel fk = kf[2 * (kM - 1) + (cols % 8 == 0 ? 0 : 1)];
        const size_t fsmem = kFW * batch_filter_warp_bytes(kM);
        int fper_sm = 0;
        int rc = batch_occupancy(h, (const void*)fk, kFW * 32, fsmem, fper_sm);
        if (rc)
            return rc;
        if (fper_sm < 1)
            return fail(h, B200LSAP_ERR_CUDA, "k_auction_batch_filter does not fit on an SM");
        int64_t fgrid = std::min<int64_t>((batch + 16 * kFW - 1) / (16 * kFW), (int64_t)fper_sm * h->num_sms);
        fk<<<(int)std::max<int64_t>(fgrid, 1), kFW * 32, fsmem, h->stream>>>(P);
    }
    else if (!smem_mode)
    {
        // lanes per problem: 8 (columns per lane <= 8) or 4 (columns per lane <= 16)
        static const int lanes_mode = getenv("B200LSAP_BATCH_LANES") ? atoi(getenv("B200LSAP_BATCH_LANES")) : 4;
        if (lanes_mode == 4)
        {
#define B200LSAP_STREAM4(C) k_auction_batch_stream<C, 4, 20>
            static const BatchKernel k4[16] = {B200LSAP_STREAM4(1),  B200LSAP_STREAM4(2),  B200LSAP_STREAM4(3),
                                               B200LSAP_STREAM4(4),  B200LSAP_STREAM4(5),  B200LSAP_STREAM4(6),
                                               B200LSAP_STREAM4(7),  B200LSAP_STREAM4(8),  B200LSAP_STREAM4(9),
                                               B200LSAP_STREAM4(10), B200LSAP_STREAM4(11), B200LSAP_STREAM4(12),
                                               B200LSAP_STREAM4(13), B200LSAP_STREAM4(14), B200LSAP_STREAM4(15),
                                               B200LSAP_STREAM4(16)};
            int64_t sgrid = (batch + 7) / 8;
            sgrid = std::min<int64_t>(sgrid, (int64_t)20 * h->num_sms);
            k4[(cols + 3) / 4 - 1]<<<(int)sgrid, 32, 0, h->stream>>>(P);
        }
        else
        {
#define B200LSAP_STREAM_TABLE(W)                                                                                   \
    {k_auction_batch_stream<1, 8, W>, k_auction_batch_stream<2, 8, W>, k_auction_batch_stream<3, 8, W>,             \
     k_auction_batch_stream<4, 8, W>, k_auction_batch_stream<5, 8, W>, k_auction_batch_stream<6, 8, W>,             \
     k_auction_batch_stream<7, 8, W>, k_auction_batch_stream<8, 8, W>}
            static const BatchKernel k16[kBatchColsPerLane] = B200LSAP_STREAM_TABLE(16);
            static const BatchKernel k24[kBatchColsPerLane] = B200LSAP_STREAM_TABLE(24);
            static const BatchKernel k32[kBatchColsPerLane] = B200LSAP_STREAM_TABLE(32);
            const int kW = warps_mode == 16 ? 16 : (warps_mode == 24 ? 24 : 32);
            const BatchKernel* tab = kW == 16 ? k16 : (kW == 32 ? k32 : k24);
            int64_t sgrid = (batch + kBatchPerWarp - 1) / kBatchPerWarp;
            sgrid = std::min<int64_t>(sgrid, (int64_t)kW * h->num_sms);
            tab[(cols + kBatchGroup - 1) / kBatchGroup - 1]<<<(int)sgrid, 32, 0, h->stream>>>(P);
        }
    }
    else
    {
        const int n_pad = (rows * cols + 3) & ~3;
        const size_t smem = (size_t)kBatchPerWarp * (n_pad + kBatchGroup) * sizeof(float);
        static const BatchKernel kernels[kBatchColsPerLane] = {k_auction_batch<1>, k_auction_batch<2>, k_auction_batch<3>,
                                                              k_auction_batch<4>, k_auction_batch<5>, k_auction_batch<6>,
                                                              k_auction_batch<7>, k_auction_batch<8>};
        const BatchKernel kernel = kernels[(cols + kBatchGroup - 1) / kBatchGroup - 1];
        int per_sm = 0;
        int rc = batch_occupancy(h, (const void*)kernel, 32, smem, per_sm);
        if (rc)
            return rc;
        if (per_sm < 1)
            return fail(h, B200LSAP_ERR_SHAPE, "batched problem does not fit in shared memory");
        const int64_t grid = std::min<int64_t>((batch + kBatchPerWarp - 1) / kBatchPerWarp, (int64_t)per_sm * h->num_sms);
        kernel<<<(int)grid, 32, smem, h->stream>>>(P);
    }
    h->launches++;
    CU(cudaEventRecord(h->ev[2], h->stream));
    CU(cudaMemcpyAsync(h->h_bctrl + slot, P.ctl, sizeof(BatchCtrl), cudaMemcpyDeviceToHost, h->stream));
    if (!sync)
        return B200LSAP_OK; // the caller synchronises and calls batch_collect
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaGetLastError());
    if (stats)
    {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]);
        stats->ms_solve = ms;
        stats->eps_final = P.eps;
        stats->bid_bytes = batch * ((int64_t)rows * cols * 4 + rows * 4);
        stats->launches = h->launches;
    }
    return batch_collect(h, slot, 1, stats);
}

int check_batch(b200lsap_handle* h, const void* a, int64_t batch, int rows, int cols, int layout, const void* out)
{
    if (!h)
        return B200LSAP_ERR_ARG;
    if (!a || !out)
        return fail(h, B200LSAP_ERR_ARG, "null pointer");
    if (layout != B200LSAP_COL_MAJOR && layout != B200LSAP_ROW_MAJOR)
        return fail(h, B200LSAP_ERR_ARG, "layout must be 0 or 1");
    if (batch < 0 || rows < 1 || cols < 1 || rows > cols || rows > kBatchMaxRows || cols > kBatchMaxCols)
        return fail(h, B200LSAP_ERR_SHAPE, "batched mode needs 1 <= rows <= cols, rows <= 32, cols <= 64");
    return B200LSAP_OK;
}

} // namespace

extern "C" {

int b200lsap_solve_batch_device_f32(b200lsap_handle* h, const float* a_dev, int64_t batch, int32_t rows, int32_t cols,
                                    int layout, const b200lsap_options* opts, int32_t* r2c_dev, b200lsap_stats* stats)
{
    int rc = check_batch(h, a_dev, batch, rows, cols, layout, r2c_dev);
    if (rc)
        return rc;
    if (stats)
        memset(stats, 0, sizeof *stats);
    if (batch == 0)
        return B200LSAP_OK;
    CU(cudaSetDevice(h->device));
    h->launches = 0;
    const float* src = a_dev;
    if (layout == B200LSAP_COL_MAJOR && rows > 1 && cols > 1)
    {
        const int64_t total = batch * rows * cols;
        CU(h->batch_in.ensure(total * sizeof(float)));
        k_batch_to_row_major<<<grid_for(h, total, 256 * 4), 256, 0, h->stream>>>(a_dev, h->batch_in.as<float>(), batch,
                                                                                rows, cols);
        h->launches++;
        src = h->batch_in.as<float>();
    }
    return run_batch(h, src, batch, rows, cols, opts, r2c_dev, stats);
}

} // extern "C"

namespace
{

// Host arrays in and out.  The batch crosses PCIe in chunks: the upload of chunk k + 1, the kernel on chunk k and the
// download of chunk k - 1 run on three streams, so the call is bound by the link (6144 bytes per 32x48 problem) and
// not by copy + solve + copy in a row.  enqueue / wait are split so that b200lsap_multi can drive several devices.
int batch_host_enqueue(b200lsap_handle* h, const float* a, int64_t batch, int rows, int cols, int layout,
                       const b200lsap_options* opts, int32_t* row_to_col, int& n_chunks)
{
    CU(cudaSetDevice(h->device));
    h->launches = 0;
    const int64_t per = (int64_t)rows * cols;
    const int64_t total = batch * per;
    CU(h->staging.ensure(total * sizeof(float)));
    CU(h->batch_out.ensure(batch * rows * sizeof(int32_t)));
    const bool convert = layout == B200LSAP_COL_MAJOR && rows > 1 && cols > 1;
    if (convert)
        CU(h->batch_in.ensure(total * sizeof(float)));
    n_chunks = (int)std::max<int64_t>(1, std::min<int64_t>(8, batch / 16384));
    CU(cudaEventRecord(h->ev[0], h->stream));
    CU(cudaStreamWaitEvent(h->up, h->ev[0], 0));   // earlier work of this handle's stream is done with the buffers
    CU(cudaStreamWaitEvent(h->down, h->ev[0], 0));
    for (int k = 0; k < n_chunks; k++)
    {
        const int64_t b0 = batch * k / n_chunks, nb = batch * (k + 1) / n_chunks - b0;
        float* stg = h->staging.as<float>() + b0 * per;
        CU(cudaMemcpyAsync(stg, a + b0 * per, nb * per * sizeof(float), cudaMemcpyHostToDevice, h->up));
        CU(cudaEventRecord(h->ev_up[k], h->up));
        CU(cudaStreamWaitEvent(h->stream, h->ev_up[k], 0));
        const float* src = stg;
        if (convert)
        {
            float* rm = h->batch_in.as<float>() + b0 * per;
            k_batch_to_row_major<<<grid_for(h, nb * per, 256 * 4), 256, 0, h->stream>>>(stg, rm, nb, rows, cols);
            h->launches++;
            src = rm;
        }
        int32_t* out = h->batch_out.as<int32_t>() + b0 * rows;
        const int rc = run_batch(h, src, nb, rows, cols, opts, out, nullptr, false, k);
        if (rc)
            return rc;
        CU(cudaEventRecord(h->ev_done[k], h->stream));
    }
    // downloads in a loop of their own: a copy into PAGEABLE host memory blocks the calling thread until its kernel has
    // finished, which would hold back the uploads of the later chunks if it were enqueued between them
    for (int k = 0; k < n_chunks; k++)
    {
        const int64_t b0 = batch * k / n_chunks, nb = batch * (k + 1) / n_chunks - b0;
        CU(cudaStreamWaitEvent(h->down, h->ev_done[k], 0));
        CU(cudaMemcpyAsync(row_to_col + b0 * rows, h->batch_out.as<int32_t>() + b0 * rows, nb * rows * sizeof(int32_t),
                           cudaMemcpyDeviceToHost, h->down));
    }
    CU(cudaEventRecord(h->ev[3], h->down));
    return B200LSAP_OK;
}

int batch_host_wait(b200lsap_handle* h, int64_t batch, int rows, int cols, const b200lsap_options* opts, int n_chunks,
                    b200lsap_stats* stats)
{
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaStreamSynchronize(h->down));
    CU(cudaStreamWaitEvent(h->stream, h->ev[3], 0)); // later work on the handle's stream follows the downloads
    CU(cudaGetLastError());
    if (stats)
    {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[0], h->ev[3]);
        stats->ms_solve = ms; // upload, kernels and download overlap: the whole pipeline
        b200lsap_options o;
        b200lsap_options_default(&o);
        if (opts)
            o = *opts;
        stats->eps_final = o.eps_multiplier / (double)cols;
        stats->bid_bytes = batch * ((int64_t)rows * cols * 4 + rows * 4);
        stats->launches = h->launches;
    }
    return batch_collect(h, 0, n_chunks, stats);
}

} // namespace

extern "C" {

int b200lsap_solve_batch_f32(b200lsap_handle* h, const float* a, int64_t batch, int32_t rows, int32_t cols, int layout,
                             const b200lsap_options* opts, int32_t* row_to_col, b200lsap_stats* stats)
{
    int rc = check_batch(h, a, batch, rows, cols, layout, row_to_col);
    if (rc)
        return rc;
    if (stats)
        memset(stats, 0, sizeof *stats);
    if (batch == 0)
        return B200LSAP_OK;
    int n_chunks = 0;
    rc = batch_host_enqueue(h, a, batch, rows, cols, layout, opts, row_to_col, n_chunks);
    if (rc)
    {
        cudaStreamSynchronize(h->up);
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->down);
        return rc;
    }
    return batch_host_wait(h, batch, rows, cols, opts, n_chunks, stats);
}

int b200lsap_generate_device_f32(b200lsap_handle* h, float* a_dev, int64_t batch, int64_t rows, int64_t cols,
                                 int64_t ld, int layout, uint64_t seed, uint64_t problem0)
{
    if (!h || !a_dev || batch < 1 || rows < 1 || cols < 1 || ld < (layout ? cols : rows))
        return fail(h, B200LSAP_ERR_ARG, "bad generate arguments");
    CU(cudaSetDevice(h->device));
    const int64_t total = batch * ld * (layout ? rows : cols);
    k_generate_f32<<<grid_for(h, total, 256 * 4), 256, 0, h->stream>>>(a_dev, batch, rows, cols, ld, layout, seed,
                                                                       problem0, 0, cols);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(h->stream));
    return B200LSAP_OK;
}

int b200lsap_generate_shard_device_f32(b200lsap_handle* h, float* a_dev, int64_t rows, int64_t cols_local, int64_t ld,
                                       int64_t col0, int64_t cols_global, uint64_t seed, uint64_t problem)
{
    if (!h || !a_dev || rows < 1 || cols_local < 1 || ld < cols_local || col0 < 0 || col0 + cols_local > cols_global)
        return fail(h, B200LSAP_ERR_ARG, "bad generate arguments");
    CU(cudaSetDevice(h->device));
    const int64_t total = ld * rows;
    k_generate_f32<<<grid_for(h, total, 256 * 4), 256, 0, h->stream>>>(a_dev, 1, rows, cols_local, ld,
                                                                       B200LSAP_ROW_MAJOR, seed, problem, col0,
                                                                       cols_global);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(h->stream));
    return B200LSAP_OK;
}

int b200lsap_bid_sweep_device_f32(b200lsap_handle* h, const float* a_dev, int64_t rows, int64_t cols, int64_t ld,
                                  int use_last_prices, double* ms_out, int64_t* bytes_out)
{
    int rc = check_shape(h, a_dev, rows, cols, ld, B200LSAP_ROW_MAJOR);
    if (rc)
        return rc;
    if (ld % 4 != 0 || cols % 4 != 0 || ((uintptr_t)a_dev & 15) != 0)
        return fail(h, B200LSAP_ERR_ARG, "bid sweep needs a 16-byte aligned row-major matrix with cols, ld % 4 == 0");
    CU(cudaSetDevice(h->device));
    int grid = 0;
    rc = persistent_grid<float>(h, grid);
    if (rc)
        return rc;
    const bool keep = use_last_prices && h->last_cols == cols;
    const bool keep_cand = keep && h->last_a == (const void*)a_dev && h->last_rows == rows && h->last_ld == ld;
    rc = alloc_state<float>(h, rows, cols, ld, grid * kWarpsPerBlock);
    if (rc)
        return rc;
    DenseState<float> s = bind_state<float>(h, a_dev, rows, cols, ld);
    s.dry = 1;
    k_init<float><<<grid, kThreads, 0, h->stream>>>(s, keep_cand ? 2 : keep ? 1 : 0);
    k_absmax<float><<<grid_for(h, rows * (int64_t)(s.cols4 / 4), 256 * 8), 256, 0, h->stream>>>(s.A, s.ld, s.rows, s.cols4,
                                                                                        s.ctrl);
    CU(cudaEventRecord(h->ev[1], h->stream));
    k_bid<float><<<grid, kThreads, 0, h->stream>>>(s, 0, (int)rows, 1e-5 / (double)cols);
    CU(cudaEventRecord(h->ev[2], h->stream));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaGetLastError());
    float ms = 0.f;
    cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]);
    if (ms_out)
        *ms_out = ms;
    if (bytes_out)
        *bytes_out = rows * cols * (int64_t)sizeof(float);
    return B200LSAP_OK;
}

} // extern "C"

// =============================================================================================
// Column-sharded solve: one big problem over several GPUs, one exchange per round.  The caller
// owns the exchange (an all-gather of the records between shard_bid and shard_resolve).
// =============================================================================================
namespace
{

DenseState<float> shard_state(b200lsap_handle* h, bool scan_view)
{
    const auto& sh = h->sh;
    DenseState<float> s = bind_state<float>(h, sh.A, sh.rows, scan_view ? sh.cols_local : sh.cols_global,
                                            scan_view ? sh.ld : round_up(sh.cols_global, 32));
    if (scan_view)
    {
        s.prices += sh.col0;
        s.pricesf += sh.col0;
    }
    return s;
}

int shard_fetch_ctrl(b200lsap_handle* h)
{
    CU(cudaMemcpyAsync(h->h_ctrl, h->ctrl.p, 4 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaGetLastError());
    return B200LSAP_OK;
}

} // namespace

extern "C" {

int64_t b200lsap_shard_record_bytes(void) { return (int64_t)sizeof(ShardRec); }

int b200lsap_shard_begin_f32(b200lsap_handle* h, const float* a_dev, int64_t rows, int64_t cols_local, int64_t ld,
                             int64_t col0, int64_t cols_global, const b200lsap_options* opts, double* amax_local)
{
    if (!h)
        return B200LSAP_ERR_ARG;
    if (!a_dev || !amax_local)
        return fail(h, B200LSAP_ERR_ARG, "null pointer");
    if (rows < 1 || cols_local < 1 || col0 < 0 || col0 + cols_local > cols_global || rows > cols_global ||
        cols_global > 0x7fff0000ll || ld < round_up(cols_local, 4))
        return fail(h, B200LSAP_ERR_SHAPE, b200lsap_status_string(B200LSAP_ERR_SHAPE));
    if (ld % 4 != 0 || col0 % 4 != 0 || ((uintptr_t)a_dev & 15) != 0)
        return fail(h, B200LSAP_ERR_ARG,
                    "a shard must be row-major, 16-byte aligned, with ld and col0 multiples of 4 (columns "
                    "[cols_local, cols_local rounded up to 4) of every row zero)");
    CU(cudaSetDevice(h->device));
    DensePlan plan;
    int rc = make_plan(h, rows, cols_global, opts, plan);
    if (rc)
        return rc;
    auto& sh = h->sh;
    sh.A = a_dev;
    sh.rows = rows;
    sh.cols_local = cols_local;
    sh.ld = ld;
    sh.col0 = col0;
    sh.cols_global = cols_global;
    sh.eps = plan.eps;
    sh.phase = 0;
    sh.cur = 0;
    sh.n = rows;
    sh.rounds = sh.bids = 0;
    sh.max_rounds = plan.max_rounds;
    rc = persistent_grid<float>(h, sh.grid);
    if (rc)
        return rc;
    rc = alloc_state<float>(h, rows, cols_global, round_up(cols_global, 32), sh.grid * kWarpsPerBlock);
    if (rc)
        return rc;
    h->launches = 0;
    DenseState<float> g = shard_state(h, false), l = shard_state(h, true);
    CU(cudaMemsetAsync(g.ctrl, 0, sizeof(DenseCtrl), h->stream));
    CU(cudaEventRecord(h->ev[1], h->stream));
    k_init<float><<<sh.grid, kThreads, 0, h->stream>>>(g, 0);
    k_absmax<float><<<grid_for(h, rows * (round_up(cols_local, 4) / 4), 256 * 8), 256, 0, h->stream>>>(l.A, l.ld, l.rows, l.cols4,
                                                                                      l.ctrl);
    h->launches += 2;
    CU(cudaMemcpyAsync(h->h_ctrl, g.ctrl, sizeof(DenseCtrl), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaGetLastError());
    *amax_local = h->h_ctrl->amax;
    sh.active = true;
    return B200LSAP_OK;
}

int b200lsap_shard_set_amax(b200lsap_handle* h, double amax_global)
{
    if (!h || !h->sh.active)
        return fail(h, B200LSAP_ERR_ARG, "no sharded solve in progress");
    CU(cudaSetDevice(h->device));
    h->h_ctrl->amax = amax_global;
    CU(cudaMemcpyAsync(&h->ctrl.as<DenseCtrl>()->amax, &h->h_ctrl->amax, sizeof(double), cudaMemcpyHostToDevice,
                       h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return B200LSAP_OK;
}

int b200lsap_shard_bid(b200lsap_handle* h, void* rec_dev, int64_t* n_out)
{
    if (!h || !h->sh.active)
        return fail(h, B200LSAP_ERR_ARG, "no sharded solve in progress");
    if (!rec_dev || !n_out)
        return fail(h, B200LSAP_ERR_ARG, "null pointer");
    auto& sh = h->sh;
    CU(cudaSetDevice(h->device));
    *n_out = sh.n;
    if (sh.n == 0)
        return B200LSAP_OK;
    if (sh.rounds >= sh.max_rounds)
        return fail(h, B200LSAP_ERR_MAX_ROUNDS, b200lsap_status_string(B200LSAP_ERR_MAX_ROUNDS));
    DenseState<float> l = shard_state(h, true);
    ShardView sv{static_cast<ShardRec*>(rec_dev), (int)sh.col0};
    k_shard_bid<float><<<sh.grid, kThreads, 0, h->stream>>>(l, sv, sh.cur, (int)sh.n);
    h->launches++;
    CU(cudaGetLastError());
    return B200LSAP_OK;
}

int b200lsap_shard_resolve(b200lsap_handle* h, const void* gathered_dev, int n_ranks, int64_t* n_next)
{
    if (!h || !h->sh.active)
        return fail(h, B200LSAP_ERR_ARG, "no sharded solve in progress");
    if (!gathered_dev || !n_next || n_ranks < 1)
        return fail(h, B200LSAP_ERR_ARG, "bad argument");
    auto& sh = h->sh;
    CU(cudaSetDevice(h->device));
    if (sh.n == 0)
    {
        *n_next = 0;
        return B200LSAP_OK;
    }
    DenseState<float> g = shard_state(h, false);
    const int n = (int)sh.n;
    k_shard_merge<float><<<grid_for(h, n, 256), 256, 0, h->stream>>>(g, static_cast<const ShardRec*>(gathered_dev),
                                                                   n_ranks, sh.cur, n, sh.eps[sh.phase]);
    k_resolve<float><<<grid_for(h, n, kThreads), kThreads, 0, h->stream>>>(g, sh.cur, n);
    k_shard_compact<<<1, 1024, 0, h->stream>>>(g.row2col, g.rows, g.list[sh.cur ^ 1], &g.ctrl->count[sh.cur ^ 1]);
    h->launches += 3;
    int rc = shard_fetch_ctrl(h);
    if (rc)
        return rc;
    if (h->h_ctrl->status != 0)
    {
        const int st = h->h_ctrl->status;
        sh.active = false;
        return fail(h, st == 4 ? B200LSAP_ERR_ARG : st,
                    st == 4 ? "sharded solve: the ranks' records are not in the same bidder order"
                            : b200lsap_status_string(st));
    }
    sh.rounds++;
    sh.bids += n;
    sh.cur ^= 1;
    sh.n = h->h_ctrl->count[sh.cur];
    *n_next = sh.n;
    return B200LSAP_OK;
}

int b200lsap_shard_next_phase(b200lsap_handle* h, int* more)
{
    if (!h || !h->sh.active || !more)
        return fail(h, B200LSAP_ERR_ARG, "no sharded solve in progress");
    auto& sh = h->sh;
    CU(cudaSetDevice(h->device));
    *more = 0;
    if (sh.n != 0 || sh.phase + 1 >= (int)sh.eps.size())
        return B200LSAP_OK;
    sh.phase++;
    DenseState<float> g = shard_state(h, false);
    k_reset<float><<<sh.grid, kThreads, 0, h->stream>>>(g, sh.cur);
    h->launches++;
    CU(cudaGetLastError());
    sh.n = sh.rows;
    *more = 1;
    return B200LSAP_OK;
}

int b200lsap_shard_finish(b200lsap_handle* h, int32_t* row_to_col_dev, double* prices_dev, b200lsap_stats* stats)
{
    if (!h || !h->sh.active)
        return fail(h, B200LSAP_ERR_ARG, "no sharded solve in progress");
    auto& sh = h->sh;
    CU(cudaSetDevice(h->device));
    CU(cudaEventRecord(h->ev[2], h->stream));
    if (row_to_col_dev)
        CU(cudaMemcpyAsync(row_to_col_dev, h->row2col.p, sh.rows * sizeof(int32_t), cudaMemcpyDeviceToDevice,
                           h->stream));
    if (prices_dev)
        CU(cudaMemcpyAsync(prices_dev, h->prices.p, sh.cols_global * sizeof(double), cudaMemcpyDeviceToDevice,
                           h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (stats)
    {
        memset(stats, 0, sizeof *stats);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]);
        stats->ms_solve = ms; // includes the caller's exchanges
        stats->rounds = sh.rounds;
        stats->bids = sh.bids;
        stats->phases = sh.phase + 1;
        stats->bid_bytes = sh.bids * sh.cols_local * (int64_t)sizeof(float);
        stats->eps_final = sh.eps.back();
        stats->launches = h->launches;
    }
    h->last_cols = sh.cols_global;
    sh.active = false;
    return sh.n == 0 ? B200LSAP_OK : fail(h, B200LSAP_ERR_ARG, "sharded solve finished with bidders left");
}

} // extern "C"

// =============================================================================================
// Fused column-sharded solve: one persistent cooperative kernel per rank, the per-round exchange inside the kernel
// (peer stores into every rank's mailbox + flags, shard_kernels.cuh).  The caller only wires the mailboxes once.
// =============================================================================================
extern "C" {

int b200lsap_shard_comm_create(b200lsap_handle* h, int64_t rows_cap, int n_ranks, void* ipc_handle_out)
{
    if (!h)
        return B200LSAP_ERR_ARG;
    if (rows_cap < 1 || n_ranks < 1 || n_ranks > kShardMaxRanks)
        return fail(h, B200LSAP_ERR_ARG, "shard_comm_create: need rows_cap >= 1 and 1 <= n_ranks <= 16");
    CU(cudaSetDevice(h->device));
    auto& f = h->fused;
    for (void* p : f.opened)
        cudaIpcCloseMemHandle(p);
    f.opened.clear();
    f.connected = false;
    f.box.release(); // a fresh allocation: a new IPC handle, flags and epoch start from zero
    const size_t bytes = kShardHeaderBytes + (size_t)2 * n_ranks * rows_cap * sizeof(ShardRec);
    CU(f.box.ensure(bytes));
    CU(cudaMemset(f.box.p, 0, bytes));
    CU(cudaDeviceSynchronize());
    f.cm = ShardComm{};
    f.cm.n_ranks = n_ranks;
    f.cm.rows_cap = rows_cap;
    f.cm.spin_ns = 10ull * 1000 * 1000 * 1000;
    if (const char* e = getenv("B200LSAP_SHARD_SPIN_MS"))
        f.cm.spin_ns = (unsigned long long)std::max(1, atoi(e)) * 1000 * 1000;
    if (ipc_handle_out)
    {
        static_assert(sizeof(cudaIpcMemHandle_t) == B200LSAP_IPC_HANDLE_BYTES, "CUDA IPC handle size");
        cudaIpcMemHandle_t ipc;
        CU(cudaIpcGetMemHandle(&ipc, f.box.p));
        memcpy(ipc_handle_out, &ipc, sizeof ipc);
    }
    return B200LSAP_OK;
}

void* b200lsap_shard_comm_ptr(b200lsap_handle* h) { return h ? h->fused.box.p : nullptr; }

int b200lsap_shard_comm_connect(b200lsap_handle* h, int rank, int n_ranks, const void* ipc_handles,
                                void* const* peer_ptrs, int ranks_on_this_device)
{
    if (!h)
        return B200LSAP_ERR_ARG;
    auto& f = h->fused;
    if (!f.box.p || n_ranks != f.cm.n_ranks || rank < 0 || rank >= n_ranks || ranks_on_this_device < 1)
        return fail(h, B200LSAP_ERR_ARG, "shard_comm_connect: create the mailbox first, with the same n_ranks");
    if (n_ranks > 1 && !ipc_handles && !peer_ptrs)
        return fail(h, B200LSAP_ERR_ARG, "shard_comm_connect: need the peers' IPC handles or their pointers");
    CU(cudaSetDevice(h->device));
    f.cm.rank = rank;
    f.share = ranks_on_this_device;
    for (int q = 0; q < n_ranks; q++)
    {
        if (q == rank)
        {
            f.cm.box[q] = static_cast<unsigned char*>(f.box.p);
            continue;
        }
        void* p = nullptr;
        if (peer_ptrs && peer_ptrs[q])
        {
            // same process: the peer's own allocation; make its device reachable from ours
            p = peer_ptrs[q];
            cudaPointerAttributes at;
            CU(cudaPointerGetAttributes(&at, p));
            if (at.device != h->device)
            {
                int can = 0;
                CU(cudaDeviceCanAccessPeer(&can, h->device, at.device));
                if (!can)
                    return fail(h, B200LSAP_ERR_CUDA, "shard_comm_connect: no peer access between the two devices");
                const cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                    return fail(h, B200LSAP_ERR_CUDA, "cudaDeviceEnablePeerAccess", e);
                cudaGetLastError();
            }
        }
        else
        {
            cudaIpcMemHandle_t ipc;
            memcpy(&ipc, static_cast<const unsigned char*>(ipc_handles) + (size_t)q * sizeof ipc, sizeof ipc);
            CU(cudaIpcOpenMemHandle(&p, ipc, cudaIpcMemLazyEnablePeerAccess));
            f.opened.push_back(p);
        }
        f.cm.box[q] = static_cast<unsigned char*>(p);
    }
    f.connected = true;
    return B200LSAP_OK;
}

int b200lsap_shard_reserve(b200lsap_handle* h, int64_t rows, int64_t cols_global)
{
    if (!h)
        return B200LSAP_ERR_ARG;
    if (rows < 1 || cols_global < rows || cols_global > 0x7fff0000ll)
        return fail(h, B200LSAP_ERR_SHAPE, b200lsap_status_string(B200LSAP_ERR_SHAPE));
    CU(cudaSetDevice(h->device));
    int grid = 0;
    int rc = persistent_grid<float>(h, grid);
    if (rc)
        return rc;
    return alloc_state<float>(h, rows, cols_global, round_up(cols_global, 32), grid * kWarpsPerBlock);
}

int b200lsap_solve_sharded_begin_f32(b200lsap_handle* h, const float* a_dev, int64_t rows, int64_t cols_local,
                                     int64_t ld, int64_t col0, int64_t cols_global, const b200lsap_options* opts)
{
    if (!h)
        return B200LSAP_ERR_ARG;
    auto& f = h->fused;
    if (!f.connected)
        return fail(h, B200LSAP_ERR_ARG, "sharded solve: the mailboxes are not connected (b200lsap_shard_comm_connect)");
    if (!a_dev)
        return fail(h, B200LSAP_ERR_ARG, "null pointer");
    if (rows < 1 || cols_local < 1 || col0 < 0 || col0 + cols_local > cols_global || rows > cols_global ||
        cols_global > 0x7fff0000ll || ld < round_up(cols_local, 4))
        return fail(h, B200LSAP_ERR_SHAPE, b200lsap_status_string(B200LSAP_ERR_SHAPE));
    if (rows > f.cm.rows_cap)
        return fail(h, B200LSAP_ERR_ARG, "sharded solve: more rows than the mailboxes were created for");
    if (ld % 4 != 0 || col0 % 4 != 0 || ((uintptr_t)a_dev & 15) != 0)
        return fail(h, B200LSAP_ERR_ARG,
                    "a shard must be row-major, 16-byte aligned, with ld and col0 multiples of 4 (columns "
                    "[cols_local, cols_local rounded up to 4) of every row zero)");
    CU(cudaSetDevice(h->device));
    DensePlan plan;
    int rc = make_plan(h, rows, cols_global, opts, plan);
    if (rc)
        return rc;
    int grid = 0;
    rc = persistent_grid<float>(h, grid);
    if (rc)
        return rc;
    grid = std::max(1, grid / f.share);
    rc = alloc_state<float>(h, rows, cols_global, round_up(cols_global, 32), grid * kWarpsPerBlock);
    if (rc)
        return rc;
    h->launches = 0;
    auto& sh = h->sh; // shard_state's inputs
    sh.A = a_dev;
    sh.rows = rows;
    sh.cols_local = cols_local;
    sh.ld = ld;
    sh.col0 = col0;
    sh.cols_global = cols_global;
    DenseState<float> g = shard_state(h, false), l = shard_state(h, true);
    DenseCtrl* hc = h->h_ctrl;
    memset(hc, 0, sizeof *hc);
    hc->count[0] = (int)rows;
    hc->n_phases = (int)plan.eps.size();
    hc->max_rounds = plan.max_rounds;
    for (size_t k = 0; k < plan.eps.size(); k++)
        hc->eps[k] = plan.eps[k];
    hc->eps_final = plan.eps.back();
    CU(cudaMemcpyAsync(g.ctrl, hc, sizeof *hc, cudaMemcpyHostToDevice, h->stream));
    CU(cudaEventRecord(h->ev[1], h->stream));
    k_init<float><<<grid, kThreads, 0, h->stream>>>(g, 0);
    // (no pass over the matrix for max |a|: the kernel takes it from its first sweep)
    int c0 = (int)col0;
    // rounds whose rows amount to at most this many bytes per rank are played by one CTA (B200LSAP_SHARD_SMALL_KB)
    static const long long small_env = getenv("B200LSAP_SHARD_SMALL_KB") ? atoll(getenv("B200LSAP_SHARD_SMALL_KB")) : -1;
    long long small_bytes = small_env >= 0 ? small_env * 1024 : 1024 * 1024;
    void* args[] = {(void*)&l, (void*)&g, (void*)&f.cm, (void*)&c0, (void*)&small_bytes};
    CU(cudaLaunchCooperativeKernel((void*)k_shard_persistent<float>, dim3(grid), dim3(kThreads), args, 0, h->stream));
    h->launches += 2;
    CU(cudaEventRecord(h->ev[2], h->stream));
    CU(cudaMemcpyAsync(h->h_ctrl, g.ctrl, sizeof(DenseCtrl), cudaMemcpyDeviceToHost, h->stream));
    f.running = true;
    f.rows = rows;
    f.cols_global = cols_global;
    f.cols_local = cols_local;
    f.eps_final = plan.eps.back();
    f.phases = (int)plan.eps.size();
    return B200LSAP_OK;
}

int b200lsap_solve_sharded_end(b200lsap_handle* h, int32_t* row_to_col_dev, double* prices_dev, b200lsap_stats* stats)
{
    if (!h || !h->fused.running)
        return fail(h, B200LSAP_ERR_ARG, "no fused sharded solve in flight");
    auto& f = h->fused;
    CU(cudaSetDevice(h->device));
    f.running = false;
    if (row_to_col_dev)
        CU(cudaMemcpyAsync(row_to_col_dev, h->row2col.p, f.rows * sizeof(int32_t), cudaMemcpyDeviceToDevice, h->stream));
    if (prices_dev)
        CU(cudaMemcpyAsync(prices_dev, h->prices.p, f.cols_global * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaGetLastError());
    if (stats)
    {
        memset(stats, 0, sizeof *stats);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]);
        stats->ms_solve = ms;
        stats->rounds = h->h_ctrl->rounds;
        stats->bids = h->h_ctrl->bids;
        stats->phases = h->h_ctrl->phase + 1;
        stats->bid_bytes = h->h_ctrl->bids * f.cols_local * (int64_t)sizeof(float);
        stats->eps_final = f.eps_final;
        stats->launches = h->launches;
    }
    h->last_cols = f.cols_global;
    const int st = h->h_ctrl->status;
    if (st != 0)
    {
        if (st == 4)
        {
            f.connected = false; // the flag counters of the ranks may have diverged: wire the mailboxes again
            return fail(h, B200LSAP_ERR_CUDA, "sharded solve: a peer rank did not arrive at an exchange (spin limit)");
        }
        return fail(h, st, b200lsap_status_string(st));
    }
    return B200LSAP_OK;
}

int b200lsap_solve_sharded_device_f32(b200lsap_handle* h, const float* a_dev, int64_t rows, int64_t cols_local,
                                      int64_t ld, int64_t col0, int64_t cols_global, const b200lsap_options* opts,
                                      int32_t* row_to_col_dev, double* prices_dev, b200lsap_stats* stats)
{
    const int rc = b200lsap_solve_sharded_begin_f32(h, a_dev, rows, cols_local, ld, col0, cols_global, opts);
    if (rc)
        return rc;
    return b200lsap_solve_sharded_end(h, row_to_col_dev, prices_dev, stats);
}

} // extern "C"

// =============================================================================================
// Several GPUs from one host thread ("the handle owns device buffers, streams and
// communicators"): b200lsap_multi owns one handle per device, enables peer access and wires the mailboxes of the
// fused sharded solve, so a C++ caller of AuctionB200.h reaches the column-sharded and the batch-split mode without
// torch.distributed or NCCL.
// =============================================================================================
struct b200lsap_multi
{
    std::vector<b200lsap_handle*> hs;
    std::vector<int> devices;
    int64_t rows_cap = 0;
    std::string err;
};

namespace
{

int mfail(b200lsap_multi* m, int status, const std::string& what)
{
    if (m)
        m->err = what;
    return status;
}

// (col0, cols_local) per rank: AuctionCommon.h:302-320 (cumulative interval ends, remainder over the first intervals)
// with boundaries on multiples of 128 columns (512 bytes), or of 4 when there are too few columns for that.
bool multi_split(int64_t cols, int n, std::vector<std::pair<int64_t, int64_t>>& out)
{
    int64_t align = 128, blocks = (cols + align - 1) / align;
    if (blocks < n)
    {
        align = 4;
        blocks = (cols + 3) / 4;
        if (blocks < n)
            return false;
    }
    out.clear();
    int64_t c = 0;
    for (int r = 0; r < n; r++)
    {
        const int64_t nb = blocks / n + (r < blocks % n ? 1 : 0);
        const int64_t c1 = std::min(cols, c + nb * align);
        out.push_back({c, c1 - c});
        c = c1;
    }
    return c == cols;
}

int multi_wire(b200lsap_multi* m, int64_t rows)
{
    if (rows <= m->rows_cap)
        return B200LSAP_OK;
    const int n = (int)m->hs.size();
    std::vector<void*> ptrs(n);
    for (int r = 0; r < n; r++)
    {
        const int rc = b200lsap_shard_comm_create(m->hs[r], rows, n, nullptr);
        if (rc)
            return mfail(m, rc, b200lsap_last_error(m->hs[r]));
        ptrs[r] = b200lsap_shard_comm_ptr(m->hs[r]);
    }
    for (int r = 0; r < n; r++)
    {
        const int share = (int)std::count(m->devices.begin(), m->devices.end(), m->devices[r]);
        const int rc = b200lsap_shard_comm_connect(m->hs[r], r, n, nullptr, ptrs.data(), share);
        if (rc)
            return mfail(m, rc, b200lsap_last_error(m->hs[r]));
    }
    m->rows_cap = rows;
    return B200LSAP_OK;
}

} // namespace

extern "C" {

int b200lsap_create_multi(b200lsap_multi** out, const int* devices, int ndev)
{
    if (!out || !devices || ndev < 1 || ndev > kShardMaxRanks)
        return B200LSAP_ERR_ARG;
    *out = nullptr;
    b200lsap_multi* m = new b200lsap_multi();
    for (int r = 0; r < ndev; r++)
    {
        b200lsap_handle* h = nullptr;
        const int rc = b200lsap_create(&h, devices[r]);
        if (rc)
        {
            for (auto* x : m->hs)
                b200lsap_destroy(x);
            delete m;
            return rc;
        }
        m->hs.push_back(h);
        m->devices.push_back(h->device);
    }
    *out = m;
    return B200LSAP_OK;
}

int b200lsap_destroy_multi(b200lsap_multi* m)
{
    if (!m)
        return B200LSAP_OK;
    for (auto* h : m->hs)
        b200lsap_destroy(h);
    delete m;
    return B200LSAP_OK;
}

const char* b200lsap_multi_last_error(const b200lsap_multi* m) { return m ? m->err.c_str() : "null handle"; }

int b200lsap_multi_devices(const b200lsap_multi* m) { return m ? (int)m->hs.size() : 0; }

int b200lsap_multi_solve_f32(b200lsap_multi* m, const float* a, int64_t rows, int64_t cols, int64_t ld, int layout,
                             const b200lsap_options* opts, int64_t* row_to_col, double* prices_out, b200lsap_stats* stats)
{
    if (!m)
        return B200LSAP_ERR_ARG;
    if (!a || !row_to_col)
        return mfail(m, B200LSAP_ERR_ARG, "null pointer");
    if (layout != B200LSAP_COL_MAJOR && layout != B200LSAP_ROW_MAJOR)
        return mfail(m, B200LSAP_ERR_ARG, "layout must be 0 or 1");
    if (rows < 1 || cols < 1 || rows > cols || ld < (layout ? cols : rows))
        return mfail(m, B200LSAP_ERR_SHAPE, b200lsap_status_string(B200LSAP_ERR_SHAPE));
    const int n = (int)m->hs.size();
    std::vector<std::pair<int64_t, int64_t>> parts;
    if (!multi_split(cols, n, parts))
        return mfail(m, B200LSAP_ERR_SHAPE, "too few columns to split over the devices in multiples of 4");
    int rc = multi_wire(m, rows);
    if (rc)
        return rc;
    if (stats)
        memset(stats, 0, sizeof *stats);
    // upload every rank's column interval (Eigen's column-major: a contiguous block, transposed on the device)
    std::vector<int64_t> lds(n);
    for (int r = 0; r < n; r++)
    {
        b200lsap_handle* h = m->hs[r];
        const int64_t c0 = parts[r].first, cl = parts[r].second, ldl = round_up(cl, 4);
        lds[r] = ldl;
        if (cudaSetDevice(h->device) != cudaSuccess || h->a.ensure((size_t)rows * ldl * sizeof(float)) != cudaSuccess)
            return mfail(m, B200LSAP_ERR_CUDA, "device buffer for a shard");
        cudaError_t e;
        if (layout == B200LSAP_COL_MAJOR)
        {
            e = h->staging.ensure((size_t)cl * rows * sizeof(float));
            if (e == cudaSuccess)
                e = cudaMemcpy2DAsync(h->staging.p, rows * sizeof(float), a + c0 * ld, ld * sizeof(float), rows * sizeof(float),
                                      cl, cudaMemcpyHostToDevice, h->stream);
            if (e == cudaSuccess)
                k_transpose_pad<float><<<grid_for(h, rows * ldl, 1024), 256, 0, h->stream>>>(h->staging.as<float>(), rows,
                                                                                           h->a.as<float>(), ldl, (int)rows, (int)cl);
        }
        else
        {
            e = cudaMemsetAsync(h->a.p, 0, (size_t)rows * ldl * sizeof(float), h->stream);
            if (e == cudaSuccess)
                e = cudaMemcpy2DAsync(h->a.p, ldl * sizeof(float), a + c0, ld * sizeof(float), cl * sizeof(float), rows,
                                      cudaMemcpyHostToDevice, h->stream);
        }
        if (e != cudaSuccess)
            return mfail(m, B200LSAP_ERR_CUDA, std::string("shard upload: ") + cudaGetErrorString(e));
        rc = b200lsap_shard_reserve(h, rows, cols); // before any rank's kernel is waiting for its peers
        if (rc)
            return mfail(m, rc, b200lsap_last_error(h));
    }
    for (int r = 0; r < n; r++)
    {
        rc = b200lsap_solve_sharded_begin_f32(m->hs[r], m->hs[r]->a.as<float>(), rows, parts[r].second, lds[r],
                                              parts[r].first, cols, opts);
        if (rc)
        {
            // ranks already started give up at the spin limit; collect them so that the handles stay usable
            for (int q = 0; q < r; q++)
                b200lsap_solve_sharded_end(m->hs[q], nullptr, nullptr, nullptr);
            m->rows_cap = 0;
            return mfail(m, rc, b200lsap_last_error(m->hs[r]));
        }
    }
    int first_rc = B200LSAP_OK;
    for (int r = 0; r < n; r++)
    {
        b200lsap_stats st;
        rc = b200lsap_solve_sharded_end(m->hs[r], nullptr, nullptr, &st);
        if (rc && !first_rc)
        {
            first_rc = rc;
            m->err = b200lsap_last_error(m->hs[r]);
        }
        if (r == 0 && stats)
            *stats = st;
        else if (stats)
            stats->ms_solve = std::max(stats->ms_solve, st.ms_solve);
    }
    if (first_rc)
    {
        if (first_rc == B200LSAP_ERR_CUDA)
            m->rows_cap = 0; // wire the mailboxes again next time
        return first_rc;
    }
    // every rank holds the same assignment and prices: rank 0's replica goes back to the caller
    b200lsap_handle* h = m->hs[0];
    if (cudaSetDevice(h->device) != cudaSuccess || h->out64.ensure(rows * sizeof(long long)) != cudaSuccess)
        return mfail(m, B200LSAP_ERR_CUDA, "result buffer");
    k_widen_i32<<<grid_for(h, rows, 256), 256, 0, h->stream>>>(h->row2col.as<int>(), h->out64.as<long long>(), rows);
    cudaMemcpyAsync(row_to_col, h->out64.p, rows * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream);
    if (prices_out)
        cudaMemcpyAsync(prices_out, h->prices.p, cols * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    if (cudaStreamSynchronize(h->stream) != cudaSuccess)
        return mfail(m, B200LSAP_ERR_CUDA, "result download");
    return B200LSAP_OK;
}

int b200lsap_multi_solve_batch_f32(b200lsap_multi* m, const float* a, int64_t batch, int32_t rows, int32_t cols,
                                   int layout, const b200lsap_options* opts, int32_t* row_to_col, b200lsap_stats* stats)
{
    if (!m)
        return B200LSAP_ERR_ARG;
    const int n = (int)m->hs.size();
    int rc = check_batch(m->hs[0], a, batch, rows, cols, layout, row_to_col);
    if (rc)
        return mfail(m, rc, b200lsap_last_error(m->hs[0]));
    if (stats)
        memset(stats, 0, sizeof *stats);
    if (batch == 0)
        return B200LSAP_OK;
    // contiguous slices of the batch, one per device; every device runs its own upload / kernel / download pipeline
    std::vector<int> chunks(n, 0);
    const int64_t per = (int64_t)rows * cols;
    int first_rc = B200LSAP_OK;
    for (int r = 0; r < n && !first_rc; r++)
    {
        const int64_t b0 = batch * r / n, nb = batch * (r + 1) / n - b0;
        if (nb == 0)
            continue;
        first_rc = batch_host_enqueue(m->hs[r], a + b0 * per, nb, rows, cols, layout, opts, row_to_col + b0 * rows, chunks[r]);
        if (first_rc)
            m->err = b200lsap_last_error(m->hs[r]);
    }
    for (int r = 0; r < n; r++)
    {
        const int64_t b0 = batch * r / n, nb = batch * (r + 1) / n - b0;
        if (nb == 0 || chunks[r] == 0)
            continue;
        b200lsap_stats st;
        memset(&st, 0, sizeof st);
        rc = batch_host_wait(m->hs[r], nb, rows, cols, opts, chunks[r], &st);
        if (rc && !first_rc)
        {
            first_rc = rc;
            m->err = b200lsap_last_error(m->hs[r]);
        }
        if (stats)
        {
            stats->ms_solve = std::max(stats->ms_solve, st.ms_solve);
            stats->bid_bytes += st.bid_bytes;
            stats->launches += st.launches;
            stats->batch_takeovers += st.batch_takeovers;
            stats->batch_max_bids = std::max(stats->batch_max_bids, st.batch_max_bids);
            stats->eps_final = st.eps_final;
        }
    }
    return first_rc;
}

} // extern "C"

// =============================================================================================
// Allocation-free batched solve (no counterpart in the reference).
// For callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid
// it.  Everything the batched kernel needs besides the caller's input and output lives in ONE file-scope device
// array: its control block (hand-out counter, status, statistics) and, for column-major problems, the staging area
// of the re-layout pass.  Every GPU that loads the library holds these B200LSAP_NOALLOC_SCRATCH_BYTES until exit.
// In host code the array's NAME is a host-side stand-in, not a device address: the address is taken once per device,
// with that device current, by cudaGetSymbolAddress, and only that pointer is handed to kernels and memsets.
// =============================================================================================
namespace
{

constexpr size_t kNoAllocCtrlBytes = 256; // BatchCtrl at the front, staging behind it (256-byte aligned)
static_assert(sizeof(BatchCtrl) <= kNoAllocCtrlBytes, "control block does not fit its slot");
static_assert(B200LSAP_NOALLOC_SCRATCH_BYTES == kNoAllocCtrlBytes + (size_t)B200LSAP_NOALLOC_MAX_ELEMS * sizeof(float),
              "header and library disagree on the scratch size");

static __device__ __align__(256) unsigned char g_noalloc_scratch[B200LSAP_NOALLOC_SCRATCH_BYTES];

constexpr int kNoAllocMaxDevices = 64;
struct NoAllocDevice
{
    bool ready = false;
    unsigned char* scratch = nullptr; // device address of g_noalloc_scratch on this device
    int num_sms = 0;
    int per_sm_filter[16] = {};       // resident CTAs per SM of the filtered kernels, 0: not asked yet
};
NoAllocDevice g_noalloc_dev[kNoAllocMaxDevices];
char g_noalloc_err[256] = "";

int noalloc_fail(int status, const char* what, cudaError_t e = cudaSuccess)
{
    if (e != cudaSuccess)
        snprintf(g_noalloc_err, sizeof g_noalloc_err, "%s: %s", what, cudaGetErrorString(e));
    else
        snprintf(g_noalloc_err, sizeof g_noalloc_err, "%s", what);
    return status;
}

#define NA(call)                                                                                                   \
    do                                                                                                             \
    {                                                                                                              \
        cudaError_t e_ = (call);                                                                                   \
        if (e_ != cudaSuccess)                                                                                     \
            return noalloc_fail(B200LSAP_ERR_CUDA, #call, e_);                                                     \
    } while (0)

} // namespace

extern "C" {

const char* b200lsap_noalloc_last_error(void) { return g_noalloc_err; }

int b200lsap_solve_batch_noalloc_f32(int device, void* cuda_stream, const float* a_dev, int64_t batch, int32_t rows,
                                     int32_t cols, int layout, const b200lsap_options* opts, int32_t* row_to_col_dev,
                                     int32_t* status_dev)
{
    if (!a_dev || !row_to_col_dev)
        return noalloc_fail(B200LSAP_ERR_ARG, "null pointer");
    if (layout != B200LSAP_COL_MAJOR && layout != B200LSAP_ROW_MAJOR)
        return noalloc_fail(B200LSAP_ERR_ARG, "layout must be 0 or 1");
    if (batch < 0 || rows < 1 || cols < 1 || rows > cols || rows > kBatchMaxRows || cols > kBatchMaxCols)
        return noalloc_fail(B200LSAP_ERR_SHAPE, "batched mode needs 1 <= rows <= cols, rows <= 32, cols <= 64");
    const bool convert = layout == B200LSAP_COL_MAJOR && rows > 1 && cols > 1;
    const int64_t total = batch * rows * cols;
    if (convert && total > (int64_t)B200LSAP_NOALLOC_MAX_ELEMS)
        return noalloc_fail(B200LSAP_ERR_SHAPE,
                            "column-major batch exceeds B200LSAP_NOALLOC_MAX_ELEMS coefficients (the static staging area); "
                            "pass row-major problems or split the batch");
    b200lsap_options o;
    b200lsap_options_default(&o);
    if (opts)
        o = *opts;
    if (!(o.eps_multiplier > 0))
        return noalloc_fail(B200LSAP_ERR_ARG, "eps_multiplier must be positive");
    if (device < 0)
        NA(cudaGetDevice(&device));
    if (device >= kNoAllocMaxDevices)
        return noalloc_fail(B200LSAP_ERR_ARG, "device index out of range");
    NA(cudaSetDevice(device));
    NoAllocDevice& d = g_noalloc_dev[device];
    if (!d.ready)
    {
        // set-up, once per device, with that device current (none of these calls allocates)
        void* p = nullptr;
        NA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        NA(cudaDeviceGetAttribute(&d.num_sms, cudaDevAttrMultiProcessorCount, device));
        d.scratch = static_cast<unsigned char*>(p);
        d.ready = true;
    }
    if (batch == 0)
        return B200LSAP_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    BatchCtrl* ctl = reinterpret_cast<BatchCtrl*>(d.scratch);
    float* staging = reinterpret_cast<float*>(d.scratch + kNoAllocCtrlBytes);
    // the array keeps the last call's contents: the control block must start at zero (counter, status, statistics)
    NA(cudaMemsetAsync(ctl, 0, sizeof(BatchCtrl), stream));
    const float* src = a_dev;
    if (convert)
    {
        const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((total + 1023) / 1024, (int64_t)d.num_sms * 8));
        k_batch_to_row_major<<<grid, 256, 0, stream>>>(a_dev, staging, batch, rows, cols);
        src = staging;
    }
    BatchParams P;
    P.a = src;
    P.row2col = row_to_col_dev;
    P.batch = batch;
    P.rows = rows;
    P.cols = cols;
    P.eps = o.eps_multiplier / (double)cols;
    P.max_bids = o.max_rounds > 0 ? (int)(o.max_rounds > 0x7fffffff ? 0x7fffffff : o.max_rounds) : 1000000;
    P.ctl = ctl;
    P.status = &ctl->status;
    P.cap = o.batch_bid_cap > 0 ? o.batch_bid_cap : (o.batch_bid_cap < 0 ? 0x7fffffff : 96);
    P.chunk = 16;
    using BatchKernel = void (*)(BatchParams);
    if (cols % 4 == 0 && ((uintptr_t)src & 15) == 0)
    {
        constexpr int kFW = 4, kFC = 4;
#define B200LSAP_NA_FILTER(M) k_auction_batch_filter<M, true, kFW, kFC>, k_auction_batch_filter<M, false, kFW, kFC>
        static const BatchKernel kf[16] = {B200LSAP_NA_FILTER(1), B200LSAP_NA_FILTER(2), B200LSAP_NA_FILTER(3),
                                           B200LSAP_NA_FILTER(4), B200LSAP_NA_FILTER(5), B200LSAP_NA_FILTER(6),
                                           B200LSAP_NA_FILTER(7), B200LSAP_NA_FILTER(8)};
        const int kM = (cols + 7) / 8;
        const int which = 2 * (kM - 1) + (cols % 8 == 0 ? 0 : 1);
        const size_t fsmem = kFW * batch_filter_warp_bytes(kM);
        if (d.per_sm_filter[which] == 0)
        {
            NA(cudaFuncSetAttribute(kf[which], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
            NA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.per_sm_filter[which], kf[which], kFW * 32, fsmem));
            if (d.per_sm_filter[which] < 1)
                return noalloc_fail(B200LSAP_ERR_CUDA, "k_auction_batch_filter does not fit on an SM");
        }
        const int64_t fgrid = std::min<int64_t>((batch + 16 * kFW - 1) / (16 * kFW), (int64_t)d.per_sm_filter[which] * d.num_sms);
        kf[which]<<<(int)std::max<int64_t>(fgrid, 1), kFW * 32, fsmem, stream>>>(P);
    }
    else
    {
#define B200LSAP_NA_STREAM4(C) k_auction_batch_stream<C, 4, 20>
        static const BatchKernel k4[16] = {B200LSAP_NA_STREAM4(1),  B200LSAP_NA_STREAM4(2),  B200LSAP_NA_STREAM4(3),
                                           B200LSAP_NA_STREAM4(4),  B200LSAP_NA_STREAM4(5),  B200LSAP_NA_STREAM4(6),
                                           B200LSAP_NA_STREAM4(7),  B200LSAP_NA_STREAM4(8),  B200LSAP_NA_STREAM4(9),
                                           B200LSAP_NA_STREAM4(10), B200LSAP_NA_STREAM4(11), B200LSAP_NA_STREAM4(12),
                                           B200LSAP_NA_STREAM4(13), B200LSAP_NA_STREAM4(14), B200LSAP_NA_STREAM4(15),
                                           B200LSAP_NA_STREAM4(16)};
        const int64_t sgrid = std::min<int64_t>((batch + 7) / 8, (int64_t)20 * d.num_sms);
        k4[(cols + 3) / 4 - 1]<<<(int)sgrid, 32, 0, stream>>>(P);
    }
    NA(cudaGetLastError());
    if (status_dev)
        NA(cudaMemcpyAsync(status_dev, &ctl->status, sizeof(int), cudaMemcpyDeviceToDevice, stream));
    return B200LSAP_OK;
}

} // extern "C"

