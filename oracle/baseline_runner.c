/*
 * baseline_runner.c -- multi-threaded CPU baseline over the oracle (TEST/BENCH
 * INFRASTRUCTURE; see nsim_oracle.c header).  Each worker owns its simulators,
 * as the reference's one-process-per-env deployment does
 * (/root/reference/nclone/run_multiple_headless.py:128-137).
 */
#define _POSIX_C_SOURCE 200809L
#include <pthread.h>
#include <time.h>
#include <stdint.h>
#include <stdlib.h>

typedef struct OSim OSim;
OSim *osim_create(void);
void osim_destroy(OSim *);
int osim_load(OSim *, const double *, int);
int osim_reset(OSim *);
void osim_tick(OSim *, int, int);
void osim_get_state(const OSim *, double *, int64_t *, double *, int64_t *);

typedef struct {
    const double *maps; const int32_t *lens; int stride; int map_lo, map_hi;
    const uint8_t *actions; int n_steps; int frame_skip; int max_frames;
    long ticks; double checksum; double loop_seconds;
} Job;

static double now_s(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

static const int ACT_H[6] = {0, -1, 1, 0, -1, 1};
static const int ACT_J[6] = {0, 0, 0, 1, 1, 1};

/* One env per map in [map_lo, map_hi): n_steps agent steps of frame_skip ticks each, the
   action of env m at step s is actions[m * n_steps + s]; Simulator.reset() on win/death or at max_frames
   (same policy as bench.py's GPU arm). */
static void *worker(void *arg) {
    Job *j = (Job *)arg;
    OSim *S = osim_create();
    long ticks = 0; double checksum = 0, loop_seconds = 0;   /* locals: the Job array is shared cache lines */
    for (int m = j->map_lo; m < j->map_hi; m++) {
        const double *md = j->maps + (size_t)m * j->stride;
        if (osim_load(S, md, j->lens[m])) continue;
        int frames = 0;
        double t0 = now_s();                 /* level load is not part of the tick path being timed */
        for (int s = 0; s < j->n_steps; s++) {
            int a = j->actions[(size_t)m * j->n_steps + s] % 6;
            double nf[31]; int64_t ni[27];
            for (int k = 0; k < j->frame_skip; k++) {
                osim_tick(S, ACT_H[a], ACT_J[a]); ticks++; frames++;
                osim_get_state(S, nf, ni, NULL, NULL);
                if (ni[0] >= 6 && ni[0] <= 8) break;      /* base_environment.py:605-609 early exit */
            }
            if ((ni[0] >= 6 && ni[0] <= 8) || frames >= j->max_frames) { osim_reset(S); frames = 0; }   /* Simulator.reset(), as the env does per episode */
            checksum += nf[0] + nf[1];
        }
        loop_seconds += now_s() - t0;
    }
    osim_destroy(S);
    j->ticks = ticks; j->checksum = checksum; j->loop_seconds = loop_seconds;
    return NULL;
}

/* returns total ticks executed */
long osim_baseline_run(const double *maps, const int32_t *lens, int n_maps, int stride,
                       const uint8_t *actions, int n_steps, int frame_skip, int max_frames,
                       int n_threads, double *checksum, double *max_loop_seconds) {
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
    long total = 0; double cs = 0, mx = 0;
    for (int t = 0; t < n_threads; t++) {
        pthread_join(th[t], NULL); total += jobs[t].ticks; cs += jobs[t].checksum;
        if (jobs[t].loop_seconds > mx) mx = jobs[t].loop_seconds;
    }
    if (checksum) *checksum = cs;
    if (max_loop_seconds) *max_loop_seconds = mx;
    free(th); free(jobs);
    return total;
}

/* ---- rollout with checkpoints (tests/test_gpu_parity.py soak slice): env e runs on map row env_map[e]; the ninja record
   (31 doubles, 27 ints: osim_get_state) after every `every`-th agent step -- before the turnover of that step -- goes to
   out_f[(s / every) * n_envs + e][31] / out_i[...][27].  Turnover = a fresh load of the level (the GPU's reset mode 0), on
   win / death or at max_frames, as tests/test_gpu_parity.py::_oracle_rollout does it. */
typedef struct {
    const double *maps; const int32_t *lens; int stride; const int32_t *env_map; int env_lo, env_hi, n_envs;
    const uint8_t *actions; int n_steps; int frame_skip; int max_frames; int every;
    double *out_f; int64_t *out_i; long ticks;
} RJob;

static void *rollout_worker(void *arg) {
    RJob *j = (RJob *)arg;
    OSim *S = osim_create();
    long ticks = 0;
    for (int e = j->env_lo; e < j->env_hi; e++) {
        const int m = j->env_map[e];
        const double *md = j->maps + (size_t)m * j->stride;
        if (osim_load(S, md, j->lens[m])) continue;
        for (int s = 0; s < j->n_steps; s++) {
            int a = j->actions[(size_t)s * j->n_envs + e] % 6;          /* actions[step][env], as the GPU takes them */
            double nf[31]; int64_t ni[27];
            for (int k = 0; k < j->frame_skip; k++) {
                osim_tick(S, ACT_H[a], ACT_J[a]); ticks++;
                osim_get_state(S, nf, ni, NULL, NULL);
                if (ni[0] >= 6 && ni[0] <= 8) break;
            }
            if ((s + 1) % j->every == 0) {
                size_t row = (size_t)(s / j->every) * j->n_envs + e;
                for (int q = 0; q < 31; q++) j->out_f[row * 31 + q] = nf[q];
                for (int q = 0; q < 27; q++) j->out_i[row * 27 + q] = ni[q];
            }
            if ((ni[0] >= 6 && ni[0] <= 8) || ni[26] >= j->max_frames) osim_load(S, md, j->lens[m]);
        }
    }
    osim_destroy(S);
    j->ticks = ticks;
    return NULL;
}

long osim_rollout_records(const double *maps, const int32_t *lens, int stride, const int32_t *env_map, int n_envs,
                          const uint8_t *actions, int n_steps, int frame_skip, int max_frames, int every, int n_threads,
                          double *out_f, int64_t *out_i) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > n_envs) n_threads = n_envs;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * n_threads);
    RJob *jobs = (RJob *)calloc(n_threads, sizeof(RJob));
    for (int t = 0; t < n_threads; t++) {
        RJob *j = &jobs[t];
        j->maps = maps; j->lens = lens; j->stride = stride; j->env_map = env_map; j->n_envs = n_envs;
        j->env_lo = (int)((long)n_envs * t / n_threads); j->env_hi = (int)((long)n_envs * (t + 1) / n_threads);
        j->actions = actions; j->n_steps = n_steps; j->frame_skip = frame_skip; j->max_frames = max_frames; j->every = every;
        j->out_f = out_f; j->out_i = out_i;
        pthread_create(&th[t], NULL, rollout_worker, j);
    }
    long total = 0;
    for (int t = 0; t < n_threads; t++) { pthread_join(th[t], NULL); total += jobs[t].ticks; }
    free(th); free(jobs);
    return total;
}
