/* Host-side cost of the "newest frame only" end-to-end path: if mg_step_host sent only the new
 * 96x96x3 frame per env (27 648 B instead of 110 592 B over PCIe), the host would have to rebuild the reference's
 * stacked observation u8[96,96,12] itself: per pixel, bytes 3..11 move to 0..8 and the new RGB goes to 9..11.
 * This times exactly that, in place, over B envs on T threads, and reports env-stacks/s and GB/s (read + write).
 * build: gcc -O3 -march=native -pthread -o host_shift host_shift.c ;  usage: host_shift [B] [threads] */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#define PIX (96 * 96)
typedef struct { uint32_t *obs; const uint8_t *frm; int e0, e1, reps; } job_t;
static void shift_env(uint32_t *o, const uint8_t *f)
{
    for (int p = 0; p < PIX; p++) {
        uint32_t w0 = o[3 * p], w1 = o[3 * p + 1], w2 = o[3 * p + 2];
        uint32_t rgb = (uint32_t)f[3 * p] | (uint32_t)f[3 * p + 1] << 8 | (uint32_t)f[3 * p + 2] << 16;
        o[3 * p] = (w0 >> 24) | (w1 << 8);
        o[3 * p + 1] = (w1 >> 24) | (w2 << 8);
        o[3 * p + 2] = (w2 >> 24) | (rgb << 8);
    }
}
static void *worker(void *a)
{
    job_t *j = (job_t *)a;
    for (int r = 0; r < j->reps; r++)
        for (int e = j->e0; e < j->e1; e++) shift_env(j->obs + (size_t)e * PIX * 3, j->frm + (size_t)e * PIX * 3);
    return 0;
}
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
int main(int argc, char **argv)
{
    int B = argc > 1 ? atoi(argv[1]) : 4096, T = argc > 2 ? atoi(argv[2]) : 16, reps = 8;
    uint32_t *obs = aligned_alloc(64, (size_t)B * PIX * 12);
    uint8_t *frm = aligned_alloc(64, (size_t)B * PIX * 3);
    memset(obs, 1, (size_t)B * PIX * 12); memset(frm, 2, (size_t)B * PIX * 3);
    pthread_t th[256]; job_t jobs[256];
    for (int pass = 0; pass < 2; pass++) {          /* pass 0 warms the pages */
        double t0 = now();
        for (int t = 0; t < T; t++) {
            jobs[t] = (job_t){ obs, frm, (int)((long long)B * t / T), (int)((long long)B * (t + 1) / T), pass ? reps : 1 };
            pthread_create(&th[t], 0, worker, &jobs[t]);
        }
        for (int t = 0; t < T; t++) pthread_join(th[t], 0);
        double dt = now() - t0;
        if (pass) {
            double per = dt / reps;
            printf("B=%d threads=%d: %.3f ms per batch of stacks, %.2f M env-stacks/s, %.1f GB/s (read %d + write %d bytes per env)\n",
                   B, T, per * 1e3, B / per / 1e6, (double)B * PIX * (12 + 3 + 12) / per / 1e9, PIX * 15, PIX * 12);
        }
    }
    return 0;
}
