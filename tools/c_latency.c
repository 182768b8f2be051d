/* qpmpc_solve timed from plain C: the C-ABI's own end-to-end latency with HOST buffers (no Python in the loop).
 * bench.py writes the problem into a file -- qpmpc_config bytes | state[24] | reference[24] | contact[4H] | best[P] |
 * timing[4] | nominal | optimize_swing (float) -- builds this client with gcc and runs it:
 *     c_latency <problem file> <solves> <warm-up>
 * Prints one line: p50_us p90_us p99_us min_us.  Every solve feeds the previous mean back in, like the control loop. */
#define _POSIX_C_SOURCE 199309L
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "qpmpc.h"

static int cmp(const void *a, const void *b) {
    const double x = *(const double *)a, y = *(const double *)b;
    return x < y ? -1 : x > y;
}

int main(int argc, char **argv) {
    if (argc < 4) return 2;
    FILE *f = fopen(argv[1], "rb");
    if (!f) return 2;
    const int K = atoi(argv[2]), W = atoi(argv[3]);
    qpmpc_config cfg;
    if (fread(&cfg, sizeof cfg, 1, f) != 1) return 2;
    qpmpc_handle *h = NULL;
    if (qpmpc_create(&cfg, &h) != QPMPC_OK) { fprintf(stderr, "qpmpc_create: %s\n", qpmpc_last_error()); return 1; }
    const int P = qpmpc_num_params(h), H = cfg.horizon;
    float state[24], ref[24], contact[4 * QPMPC_MAX_H], timing[4], tail[2];
    float *best = (float *)malloc(sizeof(float) * (size_t)P), *best_out = (float *)malloc(sizeof(float) * (size_t)P);
    if (fread(state, 4, 24, f) != 24 || fread(ref, 4, 24, f) != 24 || fread(contact, 4, (size_t)(4 * H), f) != (size_t)(4 * H) ||
        fread(best, 4, (size_t)P, f) != (size_t)P || fread(timing, 4, 4, f) != 4 || fread(tail, 4, 2, f) != 2)
        return 2;
    fclose(f);
    double *us = (double *)malloc(sizeof(double) * (size_t)K);
    qpmpc_result out;
    for (int it = 0; it < W + K; ++it) {
        struct timespec t0, t1;
        clock_gettime(CLOCK_MONOTONIC, &t0);
        const int rc = qpmpc_solve(h, state, ref, contact, best, 42, (uint64_t)it, cfg.gait_adaptive ? timing : NULL, tail[0],
                                   (int32_t)tail[1], NULL, NULL, NULL, &out, best_out, NULL, NULL);
        clock_gettime(CLOCK_MONOTONIC, &t1);
        if (rc != QPMPC_OK) { fprintf(stderr, "qpmpc_solve: %s\n", qpmpc_last_error()); return 1; }
        memcpy(best, best_out, sizeof(float) * (size_t)P);
        if (it >= W) us[it - W] = (double)(t1.tv_sec - t0.tv_sec) * 1e6 + (double)(t1.tv_nsec - t0.tv_nsec) * 1e-3;
    }
    qsort(us, (size_t)K, sizeof(double), cmp);
    printf("%.3f %.3f %.3f %.3f\n", us[K / 2], us[(K * 9) / 10], us[(K * 99) / 100], us[0]);
    qpmpc_destroy(h);
    return 0;
}
