/* The render path of the C ABI from plain C: `helm template` for a fleet of installs of ONE chart,
 * every Secret payload of the fleet encoded in GPU batches (kv_fleet_template).
 *
 *   fleet_template CHART_DIR CONFIG_FILE N    renders N installs of CHART_DIR the way README.md:60
 *       of the reference installs one (`helm install ... --set nameOverride=...,publicSshKey=...
 *       --set-file azIotEdgeConfig=config.toml`), install i named site-<i>, and writes the N
 *       `helm template` streams to stdout, each preceded by a line "#### install <i> <bytes>".
 *
 * The test (tests/test_gpu_c_example.py) compares every stream with the single-install call and
 * the Secret payload with coreutils `base64 -w0 CONFIG_FILE`.
 *
 *   gcc -std=c99 -Iinclude examples/fleet_template.c -o fleet_template \
 *       -Lkvedge_b200 -lkvedge_b200 -Wl,-rpath,$PWD/kvedge_b200
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "kvedge_b200.h"

int main(int argc, char **argv) {
    if (argc != 4) { fprintf(stderr, "usage: %s CHART_DIR CONFIG_FILE N\n", argv[0]); return 2; }
    const int n = atoi(argv[3]);
    if (n <= 0) return 2;

    /* --set-file from memory: the file is read once and handed to every install as a blob */
    FILE *f = fopen(argv[2], "rb");
    if (!f) { perror(argv[2]); return 1; }
    fseek(f, 0, SEEK_END);
    const long cfg_len = ftell(f);
    fseek(f, 0, SEEK_SET);
    uint8_t *cfg = malloc(cfg_len > 0 ? (size_t)cfg_len : 1);
    if (cfg_len > 0 && fread(cfg, 1, (size_t)cfg_len, f) != (size_t)cfg_len) { fprintf(stderr, "short read\n"); return 1; }
    fclose(f);

    kv_chart *chart = NULL;
    if (kv_chart_load(argv[1], &chart) != KV_OK) { fprintf(stderr, "load: %s\n", kv_last_error()); return 1; }

    /* one kv_overrides per install: Helm's own flag syntax for --set, a blob for --set-file */
    kv_overrides *ov = calloc((size_t)n, sizeof *ov);
    char (*sets)[96] = calloc((size_t)n, sizeof *sets);
    char (*names)[32] = calloc((size_t)n, sizeof *names);
    const char **set_ptr = calloc((size_t)n, sizeof *set_ptr), **name_ptr = calloc((size_t)n, sizeof *name_ptr);
    static const char *blob_key = "azIotEdgeConfig";
    const uint8_t *blob_data = cfg;
    size_t blob_len = (size_t)cfg_len;
    for (int i = 0; i < n; ++i) {
        snprintf(sets[i], sizeof sets[i], "nameOverride=site-%d,aziotEdgeVmEnableExternalSsh=%s", i, i % 2 ? "true" : "false");
        snprintf(names[i], sizeof names[i], "chart-%d", 1634218954 + i);
        set_ptr[i] = sets[i];
        name_ptr[i] = names[i];
        ov[i].set = &set_ptr[i];
        ov[i].n_set = 1;
        ov[i].blob_keys = &blob_key;
        ov[i].blob_data = &blob_data;
        ov[i].blob_len = &blob_len;
        ov[i].n_blob = 1;
    }

    char *out = NULL;
    int64_t *off = calloc((size_t)n + 1, sizeof *off);
    if (kv_fleet_template(chart, n, name_ptr, "default", ov, &out, off) != KV_OK) {
        fprintf(stderr, "fleet: %s\n", kv_last_error());
        return 1;
    }
    for (int i = 0; i < n; ++i) {
        printf("#### install %d %lld\n", i, (long long)(off[i + 1] - off[i]));
        fwrite(out + off[i], 1, (size_t)(off[i + 1] - off[i]), stdout);
    }
    kv_free(out);
    kv_chart_free(chart);
    kv_shutdown();
    free(off); free(ov); free(sets); free(names); free(set_ptr); free(name_ptr); free(cfg);
    return 0;
}
