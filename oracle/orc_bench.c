/*
 * orc_bench.c -- ORACLE (test infrastructure): multi-threaded CPU roll-out used
 * ONLY as the timed cpu_baseline / --impl reference leg of bench.py.
 * One env per work item, envs spread over pthreads; each env-step is the full
 * hot path of the reference: 10 physics sub-steps, allo+ego 384^2 render
 * (BaseEnv.step renders both views, base_env.py:304, 331-340), 4x downsample of
 * the three LoRes4E leaves (allo 3ch, ego 3ch, past_obs 12ch) and the score at
 * the horizon.
 */
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include "orc.h"

typedef struct {
    const mg_scene *scenes; int n_envs; int n_steps; const int32_t *actions; /* [n_steps][n_envs] */
    int tid, n_threads; double *score_sum; uint64_t *checksum; int render;
} job_t;

static void *worker(void *arg)
{
    job_t *j = (job_t *)arg;
    uint8_t *allo = malloc(MG_RES * MG_RES * 3), *ego = malloc(MG_RES * MG_RES * 3);
    uint8_t *stack384 = malloc((size_t)MG_RES * MG_RES * 12);
    uint8_t *lo_allo = malloc(MG_LORES * MG_LORES * 3), *lo_ego = malloc(MG_LORES * MG_LORES * 3);
    uint8_t *lo_stack = malloc(MG_LORES * MG_LORES * 12);
    double ssum = 0; uint64_t cks = 0;
    for (int i = j->tid; i < j->n_envs; i += j->n_threads) {
        orc_env *e = orc_create(&j->scenes[i]);
        memset(stack384, 0, (size_t)MG_RES * MG_RES * 12);
        for (int t = 0; t < j->n_steps; t++) {
            orc_env_step(e, j->actions[(size_t)t * j->n_envs + i]);
            if (e->episode_steps >= e->scene.horizon) { ssum += orc_score(e); e->episode_steps = 0; }
            if (j->render) {
                orc_render(e, 0, allo);
                orc_render(e, 1, ego);
                /* FlattenFrameStack: np.concatenate of the last four ego frames at 384^2 */
                for (size_t p = 0; p < (size_t)MG_RES * MG_RES; p++) {
                    memmove(stack384 + p * 12, stack384 + p * 12 + 3, 9);
                    memcpy(stack384 + p * 12 + 9, ego + p * 3, 3);
                }
                orc_downsample4(allo, MG_RES, MG_RES, 3, lo_allo);
                orc_downsample4(ego, MG_RES, MG_RES, 3, lo_ego);
                orc_downsample4(stack384, MG_RES, MG_RES, 12, lo_stack);
                cks += lo_stack[(t * 131) % (MG_LORES * MG_LORES * 12)] + lo_allo[7] + lo_ego[11];
            }
        }
        orc_destroy(e);
    }
    j->score_sum[j->tid] = ssum; j->checksum[j->tid] = cks;
    free(allo); free(ego); free(stack384); free(lo_allo); free(lo_ego); free(lo_stack);
    return NULL;
}

/* returns the summed scores (so the work cannot be optimised away) */
double orc_bench_rollout(const mg_scene *scenes, int n_envs, int n_steps, const int32_t *actions,
                         int n_threads, int render, uint64_t *checksum_out)
{
    if (n_threads < 1) n_threads = 1;
    pthread_t *th = malloc(sizeof(pthread_t) * n_threads);
    job_t *jobs = malloc(sizeof(job_t) * n_threads);
    double *ss = calloc(n_threads, sizeof(double));
    uint64_t *ck = calloc(n_threads, sizeof(uint64_t));
    for (int t = 0; t < n_threads; t++) {
        jobs[t] = (job_t){ scenes, n_envs, n_steps, actions, t, n_threads, ss, ck, render };
        pthread_create(&th[t], NULL, worker, &jobs[t]);
    }
    double total = 0; uint64_t c = 0;
    for (int t = 0; t < n_threads; t++) { pthread_join(th[t], NULL); total += ss[t]; c += ck[t]; }
    if (checksum_out) *checksum_out = c;
    free(th); free(jobs); free(ss); free(ck);
    return total;
}
