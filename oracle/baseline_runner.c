/*
 * baseline_runner.c -- multi-threaded CPU baseline over the oracle (TEST/BENCH
 * INFRASTRUCTURE; see nsim_oracle.c header).  Each worker owns its simulators,
 * as the reference's one-process-per-env deployment does
 * (/root/reference/nclone/run_multiple_headless.py:128-137).
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>

typedef struct OSim OSim;
OSim *osim_create(void);
void osim_destroy(OSim *);
int osim_load(OSim *, const double *, int);
void osim_tick(OSim *, int, int);
void osim_get_state(const OSim *, double *, int64_t *, double *, int64_t *);

typedef struct {
    const double *maps; const int32_t *lens; int stride; int map_lo, map_hi;
    const uint8_t *actions; int n_steps; int frame_skip; int max_frames;
    long ticks; double checksum;
} Job;

static const int ACT_H[6] = {0, -1, 1, 0, -1, 1};
static const int ACT_J[6] = {0, 0, 0, 1, 1, 1};

/* One env per map in [map_lo, map_hi): n_steps agent steps of frame_skip ticks each, the
   action of env m at step s is actions[m * n_steps + s]; reload on win/death or at max_frames
   (same policy as bench.py's GPU arm). */
static void *worker(void *arg) {
    Job *j = (Job *)arg;
    OSim *S = osim_create();
    for (int m = j->map_lo; m < j->map_hi; m++) {
        const double *md = j->maps + (size_t)m * j->stride;
        if (osim_load(S, md, j->lens[m])) continue;
        int frames = 0;
        for (int s = 0; s < j->n_steps; s++) {
            int a = j->actions[(size_t)m * j->n_steps + s] % 6;
            for (int k = 0; k < j->frame_skip; k++) { osim_tick(S, ACT_H[a], ACT_J[a]); j->ticks++; frames++; }
            double nf[31]; int64_t ni[27];
            osim_get_state(S, nf, ni, NULL, NULL);
            if (ni[0] >= 6 || frames >= j->max_frames) { osim_load(S, md, j->lens[m]); frames = 0; }
            j->checksum += nf[0] + nf[1];
        }
    }
    osim_destroy(S);
    return NULL;
}

/* returns total ticks executed */
long osim_baseline_run(const double *maps, const int32_t *lens, int n_maps, int stride,
                       const uint8_t *actions, int n_steps, int frame_skip, int max_frames,
                       int n_threads, double *checksum) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > n_maps) n_threads = n_maps;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * n_threads);
    Job *jobs = (Job *)calloc(n_threads, sizeof(Job));
    for (int t = 0; t < n_threads; t++) {
        jobs[t].maps = maps; jobs[t].lens = lens; jobs[t].stride = stride;
        jobs[t].map_lo = (int)((long)n_maps * t / n_threads);
        jobs[t].map_hi = (int)((long)n_maps * (t + 1) / n_threads);
        jobs[t].actions = actions; jobs[t].n_steps = n_steps; jobs[t].frame_skip = frame_skip; jobs[t].max_frames = max_frames;
        pthread_create(&th[t], NULL, worker, &jobs[t]);
    }
    long total = 0; double cs = 0;
    for (int t = 0; t < n_threads; t++) { pthread_join(th[t], NULL); total += jobs[t].ticks; cs += jobs[t].checksum; }
    if (checksum) *checksum = cs;
    free(th); free(jobs);
    return total;
}
