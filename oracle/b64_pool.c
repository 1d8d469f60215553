/*
 * b64_pool.c -- a persistent pthread pool around the CPU checker's encoder.
 *
 * TEST / BENCH INFRASTRUCTURE ONLY (see b64_ref.c): used by bench.py's `cpu_baseline` leg and
 * `--impl reference` arm, never by kvedge_b200.  The reference's own flow is one Helm process
 * rendering one install on one thread (README.md:56-61); a fleet of independent installs
 * (README.md:72) is what every host core can share, so the CPU arm gets all of
 * them: threads are created once, each step only wakes them, and the batch is cut by input
 * bytes (not by unit count), so that a thread with the long config.toml units does not finish
 * last.  The arithmetic is oracle_b64_encode of b64_ref.c, unchanged.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>

size_t oracle_b64_encode(const uint8_t *src, size_t n, char *dst);

typedef struct oracle_pool {
    int n_threads;
    pthread_t *tid;
    pthread_mutex_t mu;
    pthread_cond_t wake, done;
    uint64_t generation; /* bumped per job */
    int remaining;       /* workers still busy with the current job */
    int stop;
    /* the job */
    const uint8_t *in;
    const int64_t *in_off;
    char *out;
    const int64_t *out_off;
    int64_t n_units;
    int64_t *cut; /* n_threads + 1 unit indices */
} oracle_pool;

typedef struct { oracle_pool *p; int idx; } worker_arg;

static void *worker(void *argp)
{
    worker_arg *wa = (worker_arg *)argp;
    oracle_pool *p = wa->p;
    const int idx = wa->idx;
    free(wa);
    uint64_t seen = 0;
    for (;;) {
        pthread_mutex_lock(&p->mu);
        while (!p->stop && p->generation == seen)
            pthread_cond_wait(&p->wake, &p->mu);
        if (p->stop) {
            pthread_mutex_unlock(&p->mu);
            return NULL;
        }
        seen = p->generation;
        pthread_mutex_unlock(&p->mu);
        for (int64_t u = p->cut[idx]; u < p->cut[idx + 1]; u++)
            oracle_b64_encode(p->in + p->in_off[u], (size_t)(p->in_off[u + 1] - p->in_off[u]),
                              p->out + p->out_off[u]);
        pthread_mutex_lock(&p->mu);
        if (--p->remaining == 0)
            pthread_cond_signal(&p->done);
        pthread_mutex_unlock(&p->mu);
    }
}

oracle_pool *oracle_pool_create(int n_threads)
{
    if (n_threads < 1)
        n_threads = 1;
    oracle_pool *p = (oracle_pool *)calloc(1, sizeof *p);
    if (!p)
        return NULL;
    p->n_threads = n_threads;
    p->tid = (pthread_t *)calloc((size_t)n_threads, sizeof *p->tid);
    p->cut = (int64_t *)calloc((size_t)n_threads + 1, sizeof *p->cut);
    pthread_mutex_init(&p->mu, NULL);
    pthread_cond_init(&p->wake, NULL);
    pthread_cond_init(&p->done, NULL);
    for (int i = 0; i < n_threads; i++) {
        worker_arg *wa = (worker_arg *)malloc(sizeof *wa);
        wa->p = p;
        wa->idx = i;
        pthread_create(&p->tid[i], NULL, worker, wa);
    }
    return p;
}

/* Encode every unit of a packed batch on the pool's threads; returns when all are done. */
void oracle_pool_encode_batch(oracle_pool *p, const uint8_t *in, const int64_t *in_off,
                              int64_t n_units, char *out, const int64_t *out_off)
{
    /* cut by input bytes: thread t takes the units whose first byte lies in its share */
    const int64_t lo = in_off[0], total = in_off[n_units] - lo;
    int64_t u = 0;
    for (int t = 0; t < p->n_threads; t++) {
        p->cut[t] = u;
        const int64_t limit = lo + total * (t + 1) / p->n_threads;
        while (u < n_units && (in_off[u] < limit || t == p->n_threads - 1))
            u++;
    }
    p->cut[p->n_threads] = n_units;
    pthread_mutex_lock(&p->mu);
    p->in = in;
    p->in_off = in_off;
    p->out = out;
    p->out_off = out_off;
    p->n_units = n_units;
    p->remaining = p->n_threads;
    p->generation++;
    pthread_cond_broadcast(&p->wake);
    while (p->remaining > 0)
        pthread_cond_wait(&p->done, &p->mu);
    pthread_mutex_unlock(&p->mu);
}

void oracle_pool_destroy(oracle_pool *p)
{
    if (!p)
        return;
    pthread_mutex_lock(&p->mu);
    p->stop = 1;
    pthread_cond_broadcast(&p->wake);
    pthread_mutex_unlock(&p->mu);
    for (int i = 0; i < p->n_threads; i++)
        pthread_join(p->tid[i], NULL);
    pthread_mutex_destroy(&p->mu);
    pthread_cond_destroy(&p->wake);
    pthread_cond_destroy(&p->done);
    free(p->tid);
    free(p->cut);
    free(p);
}
