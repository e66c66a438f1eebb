/*
 * TEST INFRASTRUCTURE — CPU oracle for the dense auction path (see auction_oracle.h).
 * Build: make -C oracle   (gcc -O3 -march=native, no dependencies).
 */
#include "auction_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#include "../include/b200lsap_gen.h"

#define ORC_T double
#define ORC_NAME f64
#include "auction_oracle_impl.h"
#undef ORC_T
#undef ORC_NAME

#define ORC_T float
#define ORC_NAME f32
#include "auction_oracle_impl.h"
#undef ORC_T
#undef ORC_NAME

void oracle_generate_f32(float* a, int64_t rows, int64_t cols, int64_t ld, int layout, uint64_t seed,
                         uint64_t problem)
{
    const uint64_t key = b200lsap_gen_key(seed, problem);
    for (int64_t i = 0; i < rows; i++)
        for (int64_t j = 0; j < cols; j++)
        {
            const float v = b200lsap_gen_value(key, (uint64_t)(i * cols + j));
            if (layout == ORACLE_ROW_MAJOR)
                a[i * ld + j] = v;
            else
                a[i + j * ld] = v;
        }
}

void oracle_generate_f64(double* a, int64_t rows, int64_t cols, int64_t ld, int layout, uint64_t seed,
                         uint64_t problem)
{
    const uint64_t key = b200lsap_gen_key(seed, problem);
    for (int64_t i = 0; i < rows; i++)
        for (int64_t j = 0; j < cols; j++)
        {
            const double v = (double)b200lsap_gen_value(key, (uint64_t)(i * cols + j));
            if (layout == ORACLE_ROW_MAJOR)
                a[i * ld + j] = v;
            else
                a[i + j * ld] = v;
        }
}

typedef struct
{
    const float* a;
    int64_t b0, b1;
    int rows, cols, use_double;
    int32_t* row_to_col;
    int64_t failed;
} batch_job;

static void* batch_worker(void* arg)
{
    batch_job* job = (batch_job*)arg;
    const int rows = job->rows, cols = job->cols;
    const int64_t n = (int64_t)rows * cols;
    int64_t* ex = (int64_t*)malloc(sizeof(int64_t) * (size_t)cols);
    int64_t* ey = (int64_t*)malloc(sizeof(int64_t) * (size_t)cols);
    double* evd = (double*)malloc(sizeof(double) * (size_t)cols);
    float* evf = (float*)malloc(sizeof(float) * (size_t)cols);
    double* wide = (double*)malloc(sizeof(double) * (size_t)n);
    const int64_t budget = 100000 + 1000 * n;
    for (int64_t b = job->b0; b < job->b1; b++)
    {
        const float* a = job->a + b * n;
        int64_t ne = 0;
        int st;
        if (job->use_double)
        {
            for (int64_t k = 0; k < n; k++)
                wide[k] = (double)a[k];
            st = oracle_auction_f64(wide, rows, cols, cols, ORACLE_ROW_MAJOR, budget, ex, ey, evd, &ne, 0, 0,
                                    0);
        }
        else
            st = oracle_auction_f32(a, rows, cols, cols, ORACLE_ROW_MAJOR, budget, ex, ey, evf, &ne, 0, 0, 0);
        int32_t* out = job->row_to_col + b * rows;
        for (int i = 0; i < rows; i++)
            out[i] = -1;
        for (int64_t e = 0; e < ne; e++)
            out[ex[e]] = (int32_t)ey[e];
        if (st != ORC_OK)
            job->failed++;
    }
    free(ex);
    free(ey);
    free(evd);
    free(evf);
    free(wide);
    return 0;
}

int64_t oracle_auction_batch_f32(const float* a, int64_t batch, int rows, int cols, int use_double,
                                 int n_threads, int32_t* row_to_col)
{
    if (n_threads < 1)
        n_threads = 1;
    if (n_threads > 1024)
        n_threads = 1024;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)n_threads);
    batch_job* jobs = (batch_job*)malloc(sizeof(batch_job) * (size_t)n_threads);
    for (int t = 0; t < n_threads; t++)
    {
        jobs[t].a = a;
        jobs[t].b0 = batch * t / n_threads;
        jobs[t].b1 = batch * (t + 1) / n_threads;
        jobs[t].rows = rows;
        jobs[t].cols = cols;
        jobs[t].use_double = use_double;
        jobs[t].row_to_col = row_to_col;
        jobs[t].failed = 0;
        if (t > 0)
            pthread_create(&th[t], 0, batch_worker, &jobs[t]);
    }
    batch_worker(&jobs[0]);
    int64_t failed = jobs[0].failed;
    for (int t = 1; t < n_threads; t++)
    {
        pthread_join(th[t], 0);
        failed += jobs[t].failed;
    }
    free(th);
    free(jobs);
    return failed;
}
