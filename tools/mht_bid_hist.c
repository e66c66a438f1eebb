/* Developer tool (not product, not oracle): bid counts per problem of the reference's Gauss-Seidel auction
 * over the MHT shards bench.py generates (seed, problem0 = rank * batch; include/b200lsap_gen.h).
 *   gcc -O2 -fopenmp -I. tools/mht_bid_hist.c oracle/auction_oracle.c -o /tmp/mht_bid_hist -lm -lpthread
 *   /tmp/mht_bid_hist [seed=1] [batch=1048576] [shards=8] [rows=32] [cols=48]
 * Prints per shard: mean, max (and its problem id), and how many problems need more than 64/96/128/192/256/512 bids. */
#include "../oracle/auction_oracle.h"
#include <stdio.h>
#include <stdlib.h>

int main(int argc, char** argv)
{
    const uint64_t seed = argc > 1 ? strtoull(argv[1], 0, 10) : 1;
    const int64_t batch = argc > 2 ? atoll(argv[2]) : 1048576;
    const int shards = argc > 3 ? atoi(argv[3]) : 8;
    const int rows = argc > 4 ? atoi(argv[4]) : 32, cols = argc > 5 ? atoi(argv[5]) : 48;
    static const int cuts[] = {64, 96, 128, 192, 256, 512, 1024};
    for (int s = 0; s < shards; s++)
    {
        int64_t total = 0, mx = 0, mx_id = -1, over[7] = {0};
#pragma omp parallel
        {
            double* a = malloc(sizeof(double) * rows * cols);
            int64_t ex[64], ey[64], ne;
            double ev[64];
            int64_t l_total = 0, l_mx = 0, l_id = -1, l_over[7] = {0};
#pragma omp for schedule(dynamic, 1024)
            for (int64_t b = 0; b < batch; b++)
            {
                const uint64_t id = (uint64_t)s * batch + b;
                oracle_generate_f64(a, rows, cols, cols, ORACLE_ROW_MAJOR, seed, id);
                oracle_stats st;
                oracle_auction_f64(a, rows, cols, cols, ORACLE_ROW_MAJOR, -1, ex, ey, ev, &ne, 0, 0, &st);
                l_total += st.bids;
                if (st.bids > l_mx)
                    l_mx = st.bids, l_id = (int64_t)id;
                for (int c = 0; c < 7; c++)
                    l_over[c] += st.bids > cuts[c];
            }
#pragma omp critical
            {
                total += l_total;
                if (l_mx > mx)
                    mx = l_mx, mx_id = l_id;
                for (int c = 0; c < 7; c++)
                    over[c] += l_over[c];
            }
            free(a);
        }
        printf("shard %d: mean %.2f max %lld (problem %lld, local %lld) over", s, (double)total / batch, (long long)mx,
               (long long)mx_id, (long long)(mx_id - (int64_t)s * batch));
        for (int c = 0; c < 7; c++)
            printf(" >%d:%lld", cuts[c], (long long)over[c]);
        printf("\n");
        fflush(stdout);
    }
    return 0;
}
