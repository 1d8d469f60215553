/*
 * fleet_cpu.c -- the CPU twin of kv_fleet_template, for bench.py's `render_e2e` leg.
 *
 * TEST / BENCH INFRASTRUCTURE ONLY (see b64_ref.c).  Same call shape as the product's fleet
 * render (README.md:60,72: one `helm template` per install, many installs) with the GPU taken
 * out: every install is planned by the product's own host engine (kv_release_plan, handed in
 * as function pointers -- this file does not link against the product), its pending b64enc
 * sites (aziot-edge-runtime-config-secret.yaml:6, aziot-edge-vm-cloud-init-secret.yaml:6) are
 * encoded by the checker's scalar encoder (oracle_b64_encode, Go's algorithm) on the same
 * thread, and the manifests are gathered into one buffer with offsets, as kv_fleet_template
 * returns them.  All host threads share the installs.  What it answers: would a CPU-only
 * build of the same API be faster than shipping the payloads to the GPU?
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/kvedge_b200.h"

size_t oracle_b64_encode(const uint8_t *src, size_t n, char *dst);

typedef struct {
    int (*release_plan)(const kv_chart *, const char *, const char *, const kv_overrides *, kv_release **);
    int (*pending_count)(const kv_release *);
    int (*pending_input)(const kv_release *, int, const uint8_t **, size_t *);
    int (*pending_fill)(kv_release *, int, const char *, size_t);
    int (*manifest)(kv_release *, const char **, size_t *);
    void (*release_free)(kv_release *);
} oracle_kv_api;

typedef struct {
    const oracle_kv_api *api;
    const kv_chart *chart;
    int n, n_threads, idx;
    const char *const *names;
    const char *ns;
    const kv_overrides *ov;
    kv_release **rel;       /* planned and filled releases, one per install */
    size_t *len;            /* manifest length per install */
    char *out;              /* phase 2: the gathered stream */
    const int64_t *off;
    int phase, rc;
} job;

static void *work(void *argp)
{
    job *j = (job *)argp;
    char *enc = NULL;
    size_t enc_cap = 0;
    for (int i = j->idx; i < j->n; i += j->n_threads) {
        const oracle_kv_api *a = j->api;
        if (j->phase == 1) {
            kv_release *r = NULL;
            if (a->release_plan(j->chart, j->names ? j->names[i] : NULL, j->ns, j->ov ? &j->ov[i] : NULL, &r)) {
                j->rc = 1;
                break;
            }
            const int np = a->pending_count(r);
            for (int k = 0; k < np; k++) {
                const uint8_t *p;
                size_t n;
                a->pending_input(r, k, &p, &n);
                const size_t need = (n + 2) / 3 * 4;
                if (need > enc_cap) {
                    free(enc);
                    enc = (char *)malloc(need + 64);
                    enc_cap = need;
                }
                a->pending_fill(r, k, enc, oracle_b64_encode(p, n, enc));
            }
            const char *m;
            size_t mlen;
            if (a->manifest(r, &m, &mlen)) {
                j->rc = 1;
                break;
            }
            j->rel[i] = r;
            j->len[i] = mlen;
        } else {
            const char *m;
            size_t mlen;
            a->manifest(j->rel[i], &m, &mlen);
            memcpy(j->out + j->off[i], m, mlen);
            a->release_free(j->rel[i]);
        }
    }
    free(enc);
    return NULL;
}

static int run_phase(job *proto, int phase)
{
    const int T = proto->n_threads;
    pthread_t *tid = (pthread_t *)malloc(sizeof *tid * (size_t)T);
    job *jobs = (job *)malloc(sizeof *jobs * (size_t)T);
    int rc = 0;
    for (int t = 0; t < T; t++) {
        jobs[t] = *proto;
        jobs[t].idx = t;
        jobs[t].phase = phase;
        jobs[t].rc = 0;
        pthread_create(&tid[t], NULL, work, &jobs[t]);
    }
    for (int t = 0; t < T; t++) {
        pthread_join(tid[t], NULL);
        rc |= jobs[t].rc;
    }
    free(tid);
    free(jobs);
    return rc;
}

/* *out is malloc'd (free with oracle_free); out_off receives n + 1 offsets.  0 on success. */
int oracle_fleet_render_cpu(const oracle_kv_api *api, const kv_chart *chart, int n, const char *const *names,
                            const char *ns, const kv_overrides *ov, int n_threads, char **out, int64_t *out_off)
{
    if (n_threads < 1)
        n_threads = 1;
    if (n_threads > n)
        n_threads = n > 0 ? n : 1;
    job j;
    memset(&j, 0, sizeof j);
    j.api = api;
    j.chart = chart;
    j.n = n;
    j.n_threads = n_threads;
    j.names = names;
    j.ns = ns;
    j.ov = ov;
    j.rel = (kv_release **)calloc((size_t)n + 1, sizeof *j.rel);
    j.len = (size_t *)calloc((size_t)n + 1, sizeof *j.len);
    int rc = run_phase(&j, 1);
    if (!rc) {
        out_off[0] = 0;
        for (int i = 0; i < n; i++)
            out_off[i + 1] = out_off[i] + (int64_t)j.len[i];
        j.out = (char *)malloc((size_t)out_off[n] + 1);
        j.off = out_off;
        rc = run_phase(&j, 2);
        j.out[out_off[n]] = 0;
        *out = j.out;
    }
    free(j.rel);
    free(j.len);
    return rc;
}

void oracle_free(void *p) { free(p); }
