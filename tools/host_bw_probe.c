// host_bw_probe.c - how fast this box's cores assemble [n][34] float rows from two [n][17] halves (memcpy per half, k threads).
// Build + run: gcc -O3 -pthread tools/host_bw_probe.c -o /tmp/host_bw_probe && /tmp/host_bw_probe
// See tools/host_bw_probe.py for what the number decides (compact D2H + host assembly vs plain 136 B/env DMA).
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

typedef struct { float *out; const float *prev, *cur; long lo, hi; } job;
static void* work(void* p) {
    job* j = (job*)p;
    for (long i = j->lo; i < j->hi; ++i) {
        memcpy(j->out + 34 * i, j->prev + 17 * i, 68);
        memcpy(j->out + 34 * i + 17, j->cur + 17 * i, 68);
    }
    return 0;
}
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
int main(void) {
    const long n = 65536; const int shards = 8, reps = 200;
    float *out[8], *prev[8], *cur[8];
    for (int s = 0; s < shards; ++s) {
        out[s] = aligned_alloc(64, n * 136); prev[s] = aligned_alloc(64, n * 68); cur[s] = aligned_alloc(64, n * 68);
        memset(out[s], 0, n * 136); memset(prev[s], 1, n * 68); memset(cur[s], 2, n * 68);
    }
    printf("{\"n_envs\": %ld", n);
    const int ks[] = {1, 2, 4, 8, 16, 32};
    for (int q = 0; q < 6; ++q) {
        const int k = ks[q];
        pthread_t th[32]; job jb[32];
        double t0 = now();
        for (int r = 0; r < reps; ++r) {
            const int s = r % shards;
            for (int j = 0; j < k; ++j) { jb[j] = (job){out[s], prev[s], cur[s], j * n / k, (j + 1) * n / k}; pthread_create(&th[j], 0, work, &jb[j]); }
            for (int j = 0; j < k; ++j) pthread_join(th[j], 0);
        }
        const double dt = (now() - t0) / reps;
        printf(", \"%d_threads\": {\"ms_per_step\": %.4f, \"env_steps_per_s\": %.3e, \"GB_per_s_touched\": %.1f}", k, 1e3 * dt, n / dt, n * 272.0 / dt / 1e9);
    }
    printf("}\n");
    return 0;
}
