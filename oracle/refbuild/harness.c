/* ORACLE build support (test infrastructure): drives the reference's net_model_compute
 * (edge-eegnet_wolf/src/cl/net/model.c:42) over a file of trials.
 *
 *   ref_model <in.bin> <out.bin> <n_trials> [n_procs=1] [repeat=1]
 *
 * in.bin : int8 [n_trials][NET_T][NET_C_ALIGN]  (Wolf order, gen_input_header.py:75-76)
 * out.bin: int8 [n_trials][NET_N]
 * Trials are split over n_procs forked workers (the reference is not re-entrant: one arena each).
 * Prints one JSON line with the wall time of the compute loop (input already in memory).
 */
#define _GNU_SOURCE
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include "rt/rt_api.h"
#include "net.h"
#include "model.h"

void oracle_arena_reset(void);

static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: %s in.bin out.bin n_trials [n_procs] [repeat]\n", argv[0]); return 1; }
    long n = atol(argv[3]);
    int procs = argc > 4 ? atoi(argv[4]) : 1;
    int repeat = argc > 5 ? atoi(argv[5]) : 1;
    if (procs < 1) procs = 1;
    size_t trial_bytes = (size_t)NET_T * NET_C_ALIGN;
    int8_t* in = malloc(trial_bytes * n);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(in, 1, trial_bytes * n, f) != trial_bytes * n) { fprintf(stderr, "short read %s\n", argv[1]); return 1; }
    fclose(f);
    int8_t* out = mmap(NULL, (size_t)n * NET_N + 4096, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (out == MAP_FAILED) { perror("mmap"); return 1; }

    double t0 = now();
    for (int p = 0; p < procs; p++) {
        long lo = n * p / procs, hi = n * (p + 1) / procs;
        pid_t pid = procs == 1 ? 0 : fork();
        if (pid < 0) { perror("fork"); return 1; }
        if (pid == 0) {
            for (int r = 0; r < repeat; r++) {
                for (long i = lo; i < hi; i++) {
                    oracle_arena_reset();
                    int8_t* x = rt_alloc(RT_ALLOC_L2_CL_DATA, (int)trial_bytes);
                    int8_t* y = rt_alloc(RT_ALLOC_L2_CL_DATA, 16);
                    memcpy(x, in + (size_t)i * trial_bytes, trial_bytes);
                    net_model_compute(x, y);
                    memcpy(out + (size_t)i * NET_N, y, NET_N);
                }
            }
            if (procs > 1) _exit(0);
        }
    }
    int bad = 0;
    if (procs > 1) for (int p = 0; p < procs; p++) { int st; wait(&st); if (!WIFEXITED(st) || WEXITSTATUS(st)) bad = 1; }
    double t1 = now();
    if (bad) { fprintf(stderr, "worker failed\n"); return 1; }
    f = fopen(argv[2], "wb");
    fwrite(out, 1, (size_t)n * NET_N, f);
    fclose(f);
    printf("{\"trials\": %ld, \"repeat\": %d, \"procs\": %d, \"seconds\": %.6f, \"trials_per_s\": %.2f}\n",
           n, repeat, procs, t1 - t0, (double)n * repeat / (t1 - t0));
    return 0;
}
