/* The C ABI from plain C: what a Sprig/Helm maintainer's cgo shim does, without Go.
 *
 *   secret_userdata FILE...      prints, for every FILE, the line a rendered Secret carries:
 *                                "  userdata: <base64 of FILE>"   (what `{{ ... | b64enc }}` yields at
 *                                deployment/helm/templates/aziot-edge-runtime-config-secret.yaml:6)
 *
 * All files go to the GPU as ONE packed batch (kv_b64enc_batch); each result must equal
 * `base64 -w0 FILE` and decode back to FILE (BASELINE.json config 2's regression check).
 *
 *   gcc -std=c99 -Iinclude examples/secret_userdata.c -o secret_userdata \
 *       -Lkvedge_b200 -lkvedge_b200 -Wl,-rpath,$PWD/kvedge_b200
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "kvedge_b200.h"

static unsigned char *slurp(const char *path, size_t *n) {
    FILE *f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    unsigned char *p = malloc(sz > 0 ? (size_t)sz : 1);
    if (p && sz > 0 && fread(p, 1, (size_t)sz, f) != (size_t)sz) { free(p); p = NULL; }
    fclose(f);
    *n = (size_t)sz;
    return p;
}

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s FILE...\n", argv[0]); return 2; }
    const int n = argc - 1;
    int64_t *in_off = calloc((size_t)n + 1, sizeof *in_off), *out_off = calloc((size_t)n + 1, sizeof *out_off);
    unsigned char **file = calloc((size_t)n, sizeof *file);
    for (int i = 0; i < n; ++i) {
        size_t len;
        file[i] = slurp(argv[i + 1], &len);
        if (!file[i]) { fprintf(stderr, "cannot read %s\n", argv[i + 1]); return 2; }
        in_off[i + 1] = in_off[i] + (int64_t)len;
        out_off[i + 1] = out_off[i] + (int64_t)((kv_b64enc_len(len) + 15) / 16 * 16);   /* 16-byte aligned units */
    }
    /* the batch and its encodings live in page-locked memory: the copies then run at the link's rate */
    unsigned char *blob = kv_pinned_alloc((size_t)in_off[n] + 1);
    char *enc = kv_pinned_alloc((size_t)out_off[n] + 1);
    if (!blob || !enc) { fprintf(stderr, "kv_pinned_alloc: %s\n", kv_last_error()); return 1; }
    for (int i = 0; i < n; ++i) {
        memcpy(blob + in_off[i], file[i], (size_t)(in_off[i + 1] - in_off[i]));
        free(file[i]);
    }
    if (kv_b64enc_batch(blob, in_off, n, enc, out_off) != KV_OK) {
        fprintf(stderr, "kv_b64enc_batch: %s\n", kv_last_error());
        return 1;
    }
    /* the consumer side: decode every unit again and compare (the round trip of SURVEY.md 8 a6) */
    int64_t *dec_in_off = calloc((size_t)n + 1, sizeof *dec_in_off), *dec_out_off = calloc((size_t)n + 1, sizeof *dec_out_off);
    int64_t *dec_len = calloc((size_t)n, sizeof *dec_len), *status = calloc((size_t)n, sizeof *status);
    char *dense = malloc((size_t)out_off[n] + 1);
    for (int i = 0; i < n; ++i) {
        size_t el = kv_b64enc_len((size_t)(in_off[i + 1] - in_off[i]));
        memcpy(dense + dec_in_off[i], enc + out_off[i], el);
        dec_in_off[i + 1] = dec_in_off[i] + (int64_t)el;
        dec_out_off[i + 1] = dec_out_off[i] + (int64_t)((kv_b64dec_maxlen(el) + 15) / 16 * 16);
    }
    unsigned char *back = malloc((size_t)dec_out_off[n] + 1);
    if (kv_b64dec_batch(dense, dec_in_off, n, back, dec_out_off, dec_len, status) != KV_OK) {
        fprintf(stderr, "kv_b64dec_batch: %s\n", kv_last_error());
        return 1;
    }
    for (int i = 0; i < n; ++i) {
        const int64_t len = in_off[i + 1] - in_off[i];
        if (status[i] != -1 || dec_len[i] != len || memcmp(back + dec_out_off[i], blob + in_off[i], (size_t)len) != 0) {
            fprintf(stderr, "%s: round trip failed\n", argv[i + 1]);
            return 1;
        }
        printf("  userdata: %.*s\n", (int)kv_b64enc_len((size_t)len), enc + out_off[i]);
    }
    kv_pinned_free(blob);
    kv_pinned_free(enc);
    kv_shutdown();
    return 0;
}
