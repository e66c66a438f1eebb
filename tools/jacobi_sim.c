/*
 * DESIGN TOOL (not product, not oracle): CPU model of the GPU's synchronous auction schedule, used
 * to size rounds / bids / bidders-per-round for candidate schedules before spending GPU time.
 *
 *   gcc -O3 -march=native -fopenmp tools/jacobi_sim.c -o /tmp/jacobi_sim -lm
 *   /tmp/jacobi_sim rows cols seed eps_start factor tail_T topK
 *
 * Round (Jacobi): every unassigned row scans its row against the same prices, bid = a - w + eps
 * (Auction.h:250); per column the highest bid wins, lowest row among equal bids.
 * tail_T > 0: when a round has <= tail_T bidders the rows are processed one at a time (Gauss-Seidel,
 * FIFO queue) until the queue is empty or grows beyond tail_T.
 * topK > 0: models a per-row cache of the K best columns + the (K+1)-th value as a bound and
 * counts how many bids could be answered from the cache alone.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/b200lsap_gen.h"

typedef struct
{
    double v, w;
    int j;
    float a;
} top2_t;

static inline top2_t scan_row(const float* row, const double* p, int cols)
{
    top2_t t = {-1e6, -1e6, -1, 0.f};
    for (int j = 0; j < cols; j++)
    {
        const float a = row[j];
        if (a == 0.f)
            continue;
        const double d = (double)a - p[j];
        if (d > t.v)
        {
            t.w = t.v;
            t.v = d;
            t.j = j;
            t.a = a;
        }
        else if (d > t.w)
            t.w = d;
    }
    return t;
}

#define MAXK 64
typedef struct
{
    int n;            /* cached entries */
    int col[MAXK];
    double bound;     /* upper bound of a - p over every column not cached (at caching time) */
} cache_t;

static void build_cache(cache_t* c, const float* row, const double* p, int cols, int K)
{
    /* K+1 largest values */
    double val[MAXK + 1];
    int col[MAXK + 1];
    int n = 0;
    for (int j = 0; j < cols; j++)
    {
        const double d = (double)row[j] - p[j];
        if (n < K + 1 || d > val[n - 1])
        {
            int k = n < K + 1 ? n++ : n - 1;
            while (k > 0 && val[k - 1] < d)
            {
                val[k] = val[k - 1];
                col[k] = col[k - 1];
                k--;
            }
            val[k] = d;
            col[k] = j;
        }
    }
    c->n = n > K ? K : n;
    for (int k = 0; k < c->n; k++)
        c->col[k] = col[k];
    c->bound = n > K ? val[K] : -INFINITY;
}

/* policy 1 (the CTA scan of the narrow rounds): 32 column groups g = (col / 128) % 32 (the warps of a
 * 1024-thread CTA); T = the 32nd largest of the 64 values {group best, group second best}; every element
 * above T survives; the cache keeps the K best survivors, bound = T or the (K+1)-th survivor. */
static long long g_survivors = 0, g_builds1 = 0, g_overflow = 0;
static int cmp_desc(const void* a, const void* b)
{
    const double x = *(const double*)a, y = *(const double*)b;
    return x < y ? 1 : (x > y ? -1 : 0);
}
static void build_cache_stat(cache_t* c, const float* row, const double* p, int cols, int K)
{
    double hi[32], lo[32], cand[64];
    for (int g = 0; g < 32; g++)
        hi[g] = lo[g] = -INFINITY;
    for (int j = 0; j < cols; j++)
    {
        const int g = (j >> 7) & 31;
        const double d = (double)row[j] - p[j];
        if (d > hi[g])
        {
            lo[g] = hi[g];
            hi[g] = d;
        }
        else if (d > lo[g])
            lo[g] = d;
    }
    for (int g = 0; g < 32; g++)
    {
        cand[g] = hi[g];
        cand[32 + g] = lo[g];
    }
    qsort(cand, 64, sizeof(double), cmp_desc);
    const double T = cand[31];
    static __thread double sv[65536];
    static __thread int sc[65536];
    int n = 0;
    for (int j = 0; j < cols; j++)
    {
        const double d = (double)row[j] - p[j];
        if (d >= T)
        {
            sv[n] = d;
            sc[n] = j;
            n++;
        }
    }
    __sync_fetch_and_add(&g_survivors, n);
    __sync_fetch_and_add(&g_builds1, 1);
    if (n > 256)
        __sync_fetch_and_add(&g_overflow, 1);
    /* partial selection of the K+1 best */
    const int keep = n < K + 1 ? n : K + 1;
    for (int k = 0; k < keep; k++)
    {
        int m = k;
        for (int q = k + 1; q < n; q++)
            if (sv[q] > sv[m])
                m = q;
        const double tv = sv[k];
        sv[k] = sv[m];
        sv[m] = tv;
        const int tc = sc[k];
        sc[k] = sc[m];
        sc[m] = tc;
    }
    c->n = n > K ? K : n;
    for (int k = 0; k < c->n; k++)
        c->col[k] = sc[k];
    c->bound = n > K ? sv[K] : T;
}

/* returns 1 when the cache alone determines best and second best exactly */
static int cache_answers(const cache_t* c, const float* row, const double* p)
{
    double v = -INFINITY, w = -INFINITY;
    for (int k = 0; k < c->n; k++)
    {
        const double d = (double)row[c->col[k]] - p[c->col[k]];
        if (d > v)
        {
            w = v;
            v = d;
        }
        else if (d > w)
            w = d;
    }
    return w > c->bound;
}

int main(int argc, char** argv)
{
    const int rows = argc > 1 ? atoi(argv[1]) : 2048;
    const int cols = argc > 2 ? atoi(argv[2]) : rows;
    const uint64_t seed = argc > 3 ? strtoull(argv[3], 0, 10) : 1;
    const double eps_start = argc > 4 ? atof(argv[4]) : 0.25;
    const double factor = argc > 5 ? atof(argv[5]) : 5.0;
    const int tail_T = argc > 6 ? atoi(argv[6]) : 0;
    const int K = argc > 7 ? atoi(argv[7]) : 0;
    const int keep_assign = argc > 8 ? atoi(argv[8]) : 0;
    const int n_slots = argc > 9 ? atoi(argv[9]) : 0; /* GS mode: LRU of this many cached rows (0: unlimited) */
    const int early = argc > 10 ? atoi(argv[10]) : 0; /* non-final phases stop once <= early rows are unassigned */
    const int narrow_T = argc > 11 ? atoi(argv[11]) : 0; /* policy 1: caches are built and used only in rounds with <= narrow_T bidders */
    const int policy2 = argc > 12 ? atoi(argv[12]) : 0;
    long long entry_builds = 0;
    long long nr_rounds[4] = {0, 0, 0, 0}, nr_bids = 0, nr_miss = 0; /* narrow rounds with 0, 1, 2-4, >4 misses */
    int* slot_row = n_slots > 0 ? malloc(n_slots * sizeof(int)) : NULL;
    long long* slot_age = n_slots > 0 ? calloc(n_slots, sizeof(long long)) : NULL;
    for (int k = 0; k < n_slots; k++)
        slot_row[k] = -1;
    long long gs_hits = 0, gs_miss = 0, lru_clock = 0;

    float* A = malloc((size_t)rows * cols * sizeof(float));
    const uint64_t key = b200lsap_gen_key(seed, 0);
#pragma omp parallel for
    for (int i = 0; i < rows; i++)
        for (int j = 0; j < cols; j++)
            A[(size_t)i * cols + j] = b200lsap_gen_value(key, (uint64_t)i * cols + j);

    double* p = calloc(cols, sizeof(double));
    int* owner = malloc(cols * sizeof(int));
    int* r2c = malloc(rows * sizeof(int));
    int* list = malloc((size_t)rows * sizeof(int));
    int* next = malloc((size_t)rows * sizeof(int));
    top2_t* bid = malloc((size_t)rows * sizeof(top2_t));
    double* colbid = malloc(cols * sizeof(double));
    int* colrow = malloc(cols * sizeof(int));
    cache_t* cache = K > 0 ? calloc(rows, sizeof(cache_t)) : NULL;
    char* has_cache = K > 0 ? calloc(rows, 1) : NULL;

    double epsv[64];
    int nph = 0;
    const double eps_final = 1e-5 / cols;
    if (rows == cols && eps_start > 0)
        for (double e = eps_start; e > eps_final * 1.0000001 && nph < 60; e /= factor)
            epsv[nph++] = e;
    epsv[nph++] = eps_final;

    long long rounds = 0, bids = 0, gs_bids = 0, gs_episodes = 0, cache_hits = 0, cache_builds = 0;
    long long hist_r[40] = {0}, hist_b[40] = {0};
    for (int j = 0; j < cols; j++)
        owner[j] = -1;
    for (int i = 0; i < rows; i++)
        r2c[i] = -1;
    for (int ph = 0; ph < nph; ph++)
    {
        const double eps = epsv[ph];
        int n = 0;
        if (!keep_assign || ph == 0)
        {
            for (int j = 0; j < cols; j++)
                owner[j] = -1;
            for (int i = 0; i < rows; i++)
            {
                r2c[i] = -1;
                list[n++] = i;
            }
        }
        else
        {
            /* keep assignments that still satisfy eps-CS for the new eps; drop the rest */
#pragma omp parallel for
            for (int i = 0; i < rows; i++)
                bid[i] = scan_row(A + (size_t)i * cols, p, cols);
            for (int i = 0; i < rows; i++)
            {
                const int j = r2c[i];
                const double own = (double)A[(size_t)i * cols + j] - p[j];
                if (own < bid[i].v - eps)
                {
                    owner[j] = -1;
                    r2c[i] = -1;
                    list[n++] = i;
                }
            }
        }
        long long ph_rounds = 0, ph_bids = 0;
        int in_narrow = 0;
        long long ph_gs = 0;
        while (n > 0)
        {
            if (early > 0 && ph + 1 < nph && n <= early)
                break;
            if (tail_T > 0 && n <= tail_T)
            {
                /* Gauss-Seidel episode, FIFO */
                gs_episodes++;
                int head = 0, tail = n;
                int* q = list; /* circular of size rows */
                int qn = n;
                while (qn > 0 && qn <= tail_T)
                {
                    const int i = q[head];
                    head = (head + 1) % rows;
                    qn--;
                    if (K > 0)
                    {
                        int present = has_cache[i];
                        int sl = -1;
                        if (n_slots > 0)
                        {
                            present = 0;
                            for (int k = 0; k < n_slots; k++)
                                if (slot_row[k] == i)
                                {
                                    present = 1;
                                    sl = k;
                                }
                            if (!present)
                            {
                                sl = 0;
                                for (int k = 1; k < n_slots; k++)
                                    if (slot_age[k] < slot_age[sl])
                                        sl = k;
                            }
                            slot_age[sl] = ++lru_clock;
                            slot_row[sl] = i;
                        }
                        if (present && has_cache[i] && cache_answers(&cache[i], A + (size_t)i * cols, p))
                        {
                            cache_hits++;
                            gs_hits++;
                        }
                        else
                        {
                            if (narrow_T > 0)
                                build_cache_stat(&cache[i], A + (size_t)i * cols, p, cols, K);
                            else
                                build_cache(&cache[i], A + (size_t)i * cols, p, cols, K);
                            has_cache[i] = 1;
                            cache_builds++;
                            gs_miss++;
                        }
                    }
                    const top2_t t = scan_row(A + (size_t)i * cols, p, cols);
                    const double b = (double)t.a - t.w + eps;
                    p[t.j] = b;
                    const int prev = owner[t.j];
                    owner[t.j] = i;
                    r2c[i] = t.j;
                    if (prev >= 0)
                    {
                        r2c[prev] = -1;
                        q[tail % rows] = prev;
                        tail = (tail + 1) % rows;
                        qn++;
                    }
                    gs_bids++;
                    ph_gs++;
                    bids++;
                    ph_bids++;
                }
                /* compact queue back to list[0..qn) */
                int m = 0;
                for (int k = 0; k < qn; k++)
                    next[m++] = q[(head + k) % rows];
                memcpy(list, next, m * sizeof(int));
                n = m;
                continue;
            }
            int b = 0;
            while ((1 << (b + 1)) <= n)
                b++;
            hist_r[b]++;
            hist_b[b] += n;
#pragma omp parallel for schedule(dynamic, 16) if (n >= 256)
            for (int k = 0; k < n; k++)
            {
                const int i = list[k];
                bid[k] = scan_row(A + (size_t)i * cols, p, cols);
            }
            if (K > 0 && narrow_T > 0)
            {
                if (n <= narrow_T)
                {
                    int miss = 0;
                    if (!in_narrow && policy2)
                    {
                        /* policy 2: entering the narrow regime of this phase, every row gets a fresh cache */
#pragma omp parallel for schedule(dynamic, 16)
                        for (int i = 0; i < rows; i++)
                        {
                            build_cache_stat(&cache[i], A + (size_t)i * cols, p, cols, K);
                            has_cache[i] = 1;
                        }
                        entry_builds += rows;
                    }
                    in_narrow = 1;
                    for (int k = 0; k < n; k++)
                    {
                        const int i = list[k];
                        if (has_cache[i] && cache_answers(&cache[i], A + (size_t)i * cols, p))
                            cache_hits++;
                        else
                        {
                            build_cache_stat(&cache[i], A + (size_t)i * cols, p, cols, K);
                            has_cache[i] = 1;
                            cache_builds++;
                            miss++;
                        }
                    }
                    nr_rounds[miss == 0 ? 0 : miss == 1 ? 1 : miss <= 4 ? 2 : 3]++;
                    nr_bids += n;
                    nr_miss += miss;
                }
            }
            else if (K > 0)
            {
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : cache_hits, cache_builds) if (n >= 256)
                for (int k = 0; k < n; k++)
                {
                    const int i = list[k];
                    if (has_cache[i] && cache_answers(&cache[i], A + (size_t)i * cols, p))
                        cache_hits++;
                    else
                    {
                        build_cache(&cache[i], A + (size_t)i * cols, p, cols, K);
                        has_cache[i] = 1;
                        cache_builds++;
                    }
                }
            }
            for (int k = 0; k < n; k++)
            {
                const int j = bid[k].j;
                colrow[j] = -1;
            }
            for (int k = 0; k < n; k++)
            {
                const int i = list[k], j = bid[k].j;
                const double bv = (double)bid[k].a - bid[k].w + eps;
                if (colrow[j] < 0 || bv > colbid[j] || (bv == colbid[j] && i < colrow[j]))
                {
                    colbid[j] = bv;
                    colrow[j] = i;
                }
            }
            int m = 0;
            for (int k = 0; k < n; k++)
            {
                const int i = list[k], j = bid[k].j;
                if (colrow[j] == i)
                {
                    p[j] = colbid[j];
                    const int prev = owner[j];
                    owner[j] = i;
                    r2c[i] = j;
                    if (prev >= 0)
                    {
                        r2c[prev] = -1;
                        next[m++] = prev;
                    }
                }
                else
                    next[m++] = i;
            }
            int* t = list;
            list = next;
            next = t;
            rounds++;
            ph_rounds++;
            bids += n;
            ph_bids += n;
            n = m;
        }
        fprintf(stderr, "phase %2d eps %.3e: rounds %lld bids %lld (gs %lld)\n", ph, eps, ph_rounds, ph_bids, ph_gs);
    }
    double obj = 0;
    for (int i = 0; i < rows; i++)
        obj += (double)A[(size_t)i * cols + r2c[i]];
    printf("rows %d cols %d seed %llu eps_start %g factor %g tail_T %d K %d keep %d\n", rows, cols,
           (unsigned long long)seed, eps_start, factor, tail_T, K, keep_assign);
    printf("jacobi rounds %lld, total bids %lld (gs bids %lld in %lld episodes), objective %.9f\n", rounds, bids,
           gs_bids, gs_episodes, obj);
    if (K > 0)
        printf("GS-mode cache (slots %d): hits %lld misses %lld\n", n_slots, gs_hits, gs_miss);
    if (K > 0)
        printf("cache: hits %lld, builds (full scans) %lld\n", cache_hits, cache_builds);
    {
        /* cost model (us): wide round 16 + bandwidth at 5 TB/s, narrow round 1.5 + 2.5 per miss, tail bid 1.2 + 2.5 per miss,
         * cache build sweep per phase = one matrix read at 2.5 TB/s */
        const long long narrow_r = nr_rounds[0] + nr_rounds[1] + nr_rounds[2] + nr_rounds[3];
        const long long wide_r = rounds - narrow_r;
        const long long wide_b = bids - gs_bids - nr_bids;
        const double row_us = cols * 4.0 / 5.0e6;
        const double cost = 16.0 * wide_r + row_us * wide_b + 1.5 * narrow_r + 2.5 * (nr_miss + gs_miss) + 1.2 * gs_bids +
                            (policy2 ? entry_builds * row_us * 2.0 : 0.0);
        printf("COST %.1f ms: phases %d wide rounds %lld (bids %lld) narrow rounds %lld (misses %lld) tail bids %lld (misses %lld)\n",
               cost / 1e3, nph, wide_r, wide_b, narrow_r, nr_miss, gs_bids, gs_miss);
    }
    if (policy2)
        printf("policy 2: %lld cache builds at narrow-regime entries\n", entry_builds);
    if (narrow_T > 0)
        printf("policy 1 (narrow <= %d): narrow rounds with 0 / 1 / 2-4 / >4 misses: %lld / %lld / %lld / %lld; bids %lld misses %lld; "
               "survivors per build %.1f, builds %lld, >256 survivors %lld\n",
               narrow_T, nr_rounds[0], nr_rounds[1], nr_rounds[2], nr_rounds[3], nr_bids, nr_miss,
               g_builds1 ? (double)g_survivors / g_builds1 : 0.0, g_builds1, g_overflow);
    printf("bidders/round histogram (jacobi rounds only):\n");
    for (int b = 0; b < 40; b++)
        if (hist_r[b])
            printf("  [%6d, %6d): rounds %8lld bids %10lld\n", 1 << b, 1 << (b + 1), hist_r[b], hist_b[b]);
    return 0;
}
