/*
 * kvedge_b200.h -- C ABI of libkvedge_b200.so
 *
 * Drop-in boundary for the ONE byte path the reference (levi106/kvedge) drives:
 *
 *     values -> Helm render -> b64enc -> Secret.data.userdata        
 *
 * The reference defines no FFI, plugin or operator interface of its own (it is a Helm
 * chart).  The entry points below are therefore shaped after what a
 * maintainer of the reference's *executor* (Helm's template function map, Sprig) would
 * bind for this path: plain pointers and sizes, caller-owned buffers, no torch types.
 * Each entry cites the reference site it serves (paths relative to /root/reference).
 *
 * All functions return KV_OK (0) or a negative KV_E* code; kv_last_error() returns the
 * message of the last failure on the calling thread.  Nothing here falls back to the
 * CPU: without a CUDA device every compute entry returns KV_ENODEVICE.
 *
 * Batch layout ("packed batch"): n_units independent byte strings in one blob, unit i
 * occupying bytes [off[i], off[i+1]) of the blob; `off` holds n_units+1 int64 entries.
 * One unit = one Secret payload of one install.  For outputs, unit i is written at
 * out_off[i] and may use up to out_off[i+1]-out_off[i] bytes (capacity); 16-byte aligned
 * output starts take the fastest store path.
 */
#ifndef KVEDGE_B200_H_
#define KVEDGE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KV_OK 0
#define KV_EINVAL (-1)    /* bad argument */
#define KV_ENODEVICE (-2) /* no CUDA device / driver: there is no CPU fallback */
#define KV_ECUDA (-3)     /* a CUDA runtime call failed */
#define KV_ENOMEM (-4)    /* scratch or staging too small / allocation failed */
#define KV_ECORRUPT (-5)  /* base64 input is malformed (decode) */
#define KV_ETEMPLATE (-6) /* chart load / template parse / execution error */

/* Library ABI version (major*10000 + minor*100 + patch). */
int kv_version(void);
/* Message of the last error raised on this thread ("" if none). */
const char *kv_last_error(void);
/* Number of visible CUDA devices (0 when there is no driver). Never fails. */
int kv_device_count(void);

/* ---- sizes ------------------------------------------------------------------------
 * Go encoding/base64 StdEncoding.EncodedLen / DecodedLen, as used by Sprig `b64enc` at
 *   deployment/helm/templates/aziot-edge-runtime-config-secret.yaml:6
 *   deployment/helm/templates/aziot-edge-vm-cloud-init-secret.yaml:6              */
size_t kv_b64enc_len(size_t n);    /* 4 * ceil(n / 3) */
size_t kv_b64dec_maxlen(size_t n); /* 3 * (n / 4): upper bound of the decoded size */

/* ---- device-resident path (inputs and outputs already in HBM) -----------------------
 * Scratch needed by the *_dev entry points for a batch of n_units units totalling
 * total_in_bytes input bytes (tile table; see DESIGN.md "data layout").             */
size_t kv_b64_scratch_bytes(int64_t n_units, int64_t total_in_bytes);

/* Sprig b64enc over a packed batch (replaces the two `| b64enc` pipeline stages at
 * aziot-edge-runtime-config-secret.yaml:6 and aziot-edge-vm-cloud-init-secret.yaml:6,
 * one unit per rendered Secret).  All pointers are device pointers; `stream` is a
 * cudaStream_t (NULL = default stream).  total_in_bytes must be >= d_in_off[n_units]
 * (it sizes the tile table; the offsets themselves are only read on the device).
 * Asynchronous: returns after enqueueing.                                          */
int kv_b64enc_batch_dev(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                        int64_t total_in_bytes, char *d_out, const int64_t *d_out_off,
                        void *d_scratch, size_t scratch_bytes, void *stream);

/* The same call split in two, for callers that re-encode new bytes under unchanged offsets
 * (same fleet, new configs): kv_b64enc_plan_dev builds the tile table in d_scratch from the
 * offsets, kv_b64enc_run_dev runs the encode kernel over it.  plan + run == batch_dev.
 * The table remembers d_out_off (short units share tiles and look their output position up
 * there), so both offset arrays must stay where they are, unchanged, while a plan is reused. */
int kv_b64enc_plan_dev(const int64_t *d_in_off, int64_t n_units, int64_t total_in_bytes,
                       const int64_t *d_out_off, void *d_scratch, size_t scratch_bytes,
                       void *stream);
int kv_b64enc_run_dev(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                      int64_t total_in_bytes, char *d_out, void *d_scratch,
                      size_t scratch_bytes, void *stream);

/* kv_b64enc_batch_dev for callers that cannot supply scratch memory.  (No counterpart in the
 * reference: this exists for the embedding side, not for the chart.)  Some hosts of this
 * library are not able to allocate device memory at all: a process whose HBM was claimed up
 * front by a serving engine or a framework's caching pool, a latency-critical loop that must
 * not call cudaMalloc/cudaFree (both may synchronise the device), or a plugin that is only
 * ever handed input and output pointers.  For them the tile table lives in a fixed array of
 * KV_STATIC_SCRATCH_BYTES inside the library's device image; this entry point allocates
 * nothing, on the host or on the device.
 *  - Limit: the batch's tile table must fit in KV_STATIC_SCRATCH_BYTES, that is
 *        512 + 256 * ceil(ceil(n_units / 1024) / 64) + 32 * (n_units + total_in_bytes / 12288)
 *    bytes (32 bytes per unit plus 32 bytes per 12 KiB of input, in practice; a little less
 *    than kv_b64_scratch_bytes() reports, which is sized for decode as well).  Larger batches
 *    are refused with KV_ENOMEM and a message saying so.  Every GPU that loads the library
 *    holds the array until the process exits.
 *  - The first call on each device looks the array's device address up and does the encode
 *    kernels' one-time launch set-up (once, with that device current).  The driver loads this
 *    library's device code and the array itself at that moment, which takes device memory
 *    even though the library allocates none: make that first call at start-up -- before
 *    another pool claims the remaining memory, and outside any stream capture.  A batch of
 *    one empty unit is enough.
 *  - All calls on one GPU share the array, which keeps whatever the previous call left in
 *    it (each call rewrites everything it reads).  Two calls whose kernels overlap on the
 *    same GPU -- different streams, different threads -- race on it: serialise them.
 * Same arguments and results as kv_b64enc_batch_dev otherwise; asynchronous on `stream`.  */
#define KV_STATIC_SCRATCH_BYTES (8u << 20)
int kv_b64enc_batch_dev_static(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                               int64_t total_in_bytes, char *d_out, const int64_t *d_out_off,
                               void *stream);

/* The *_dev entries are asynchronous and read the offsets only on the device, so they cannot
 * tell at call time whether total_in_bytes was at least d_in_off[n_units].  If it was smaller,
 * the tile table is clamped to the size it was given: memory stays intact, but units past the
 * clamp are NOT processed and the call has long returned KV_OK.  This entry reports it: it waits
 * for `stream`, reads the plan's verdict out of d_scratch and returns KV_EINVAL when the last
 * batch planned there was truncated.  (No counterpart in the reference.  The static entry keeps
 * its table to itself and cannot be asked: give it a correct total_in_bytes.)                 */
int kv_b64_batch_status_dev(const void *d_scratch, void *stream);

/* Inverse of the above: what the Kubernetes API server / KubeVirt do to
 * Secret.data.userdata before exposing it as the `userdata` file consumed at
 * deployment/helm/templates/_helper.tpl:64,73 (via aziot-edge-vm.yaml:25-28,49-51).
 * Go StdEncoding.DecodeString semantics: padding required, '\r' and '\n' ignored,
 * trailing bits not checked.  Per unit: d_out_len[i] = decoded size; d_status[i] = -1
 * on success, else the byte offset Go reports in CorruptInputError (the unit's output
 * is then unspecified).  Asynchronous.                                             */
int kv_b64dec_batch_dev(const char *d_in, const int64_t *d_in_off, int64_t n_units,
                        int64_t total_in_bytes, uint8_t *d_out, const int64_t *d_out_off,
                        int64_t *d_out_len, int64_t *d_status, void *d_scratch,
                        size_t scratch_bytes, void *stream);

/* ---- host-buffer path (what a template engine calls) --------------------------------
 * `func b64enc(string) string` of Sprig, as invoked at
 * aziot-edge-runtime-config-secret.yaml:6 / aziot-edge-vm-cloud-init-secret.yaml:6.
 * `out` must hold kv_b64enc_len(n) bytes; no NUL terminator is written.  Synchronous:
 * copies to the device, encodes there, copies back.                                 */
int kv_b64enc(const uint8_t *in, size_t n, char *out);
/* Batched form over host buffers (pinned or pageable).  Synchronous; chunks the batch
 * and overlaps H2D, kernel and D2H on three streams.                               */
int kv_b64enc_batch(const uint8_t *in, const int64_t *in_off, int64_t n_units, char *out,
                    const int64_t *out_off);
/* Host-buffer decode of one string; *out_n receives the decoded size.  Returns
 * KV_ECORRUPT and stores Go's CorruptInputError offset in *err_off on malformed input. */
int kv_b64dec(const char *in, size_t n, uint8_t *out, size_t *out_n, size_t *err_off);
int kv_b64dec_batch(const char *in, const int64_t *in_off, int64_t n_units, uint8_t *out,
                    const int64_t *out_off, int64_t *out_len, int64_t *status);

/* ---- render path: values -> templates -> (pending b64enc) -> `helm template` stream -----
 * What Helm does with the reference's chart (README.md:60:
 *   helm install --generate-name . --set publicSshKey=...,nameOverride=... --set-file azIotEdgeConfig=config.toml)
 * restricted to the Go-template / Sprig subset the chart uses
 * (deployment/helm/templates/_helper.tpl:1-75 and the six resource templates),
 * with deployment/helm/.helmignore honoured (l.25 disables the HTTP-disk DataVolume).
 * A release is first *planned*: every template is rendered on the host, but each top-level
 * `... | b64enc` site (aziot-edge-runtime-config-secret.yaml:6, aziot-edge-vm-cloud-init-secret.yaml:6)
 * is left pending; kv_release_encode_gpu then encodes all pending payloads in one GPU batch.
 * A b64enc whose result feeds another function is evaluated at once (still on the GPU).   */
typedef struct kv_chart kv_chart;
typedef struct kv_release kv_release;

/* Value overrides in Helm's own flag syntax (pkg/strvals). `blob_*` is --set-file with the
 * file content already in memory (verbatim bytes). */
typedef struct kv_overrides {
    const char *const *set;        int n_set;         /* --set k=v[,k=v]          */
    const char *const *set_string; int n_set_string;  /* --set-string k=v         */
    const char *const *set_file;   int n_set_file;    /* --set-file k=path        */
    const char *const *blob_keys;  const uint8_t *const *blob_data;
    const size_t *blob_len;        int n_blob;        /* --set-file, from memory  */
} kv_overrides;

int kv_chart_load(const char *chart_dir, kv_chart **out);
void kv_chart_free(kv_chart *chart);
/* Render one release with its b64enc sites pending (no GPU work yet). */
int kv_release_plan(const kv_chart *chart, const char *release_name, const char *ns,
                    const kv_overrides *overrides, kv_release **out);
int kv_release_pending_count(const kv_release *r);
/* Input bytes of pending site i (valid until the release is freed). */
int kv_release_pending_input(const kv_release *r, int i, const uint8_t **data, size_t *len);
/* Supply the encoding of pending site i from outside (a caller that batches many releases
 * through kv_b64enc_batch_dev itself); len must be kv_b64enc_len(input len). */
int kv_release_pending_fill(kv_release *r, int i, const char *enc, size_t len);
/* Encode every pending site on the GPU (one kv_b64enc_batch call). */
int kv_release_encode_gpu(kv_release *r);
/* `helm template` stream ("---\n# Source: <chart>/templates/<file>\n<doc>\n" per document, in
 * Helm's install order) and the NOTES text (what `helm install` prints; the reference's
 * only recorded real-Helm output is deployment/az-iot-edge-k8s-kubevirt-ascii.cast:2294).
 * Fail with KV_EINVAL while sites are still pending. Pointers stay valid until free. */
int kv_release_manifest(kv_release *r, const char **data, size_t *len);
int kv_release_notes(kv_release *r, const char **data, size_t *len);
void kv_release_free(kv_release *r);
/* One call: load, plan, encode on the GPU, frame. *out is malloc'd: release with kv_free. */
int kv_helm_template(const char *chart_dir, const char *release_name, const char *ns,
                     const kv_overrides *overrides, char **out, size_t *out_len);
/* n_releases installs of one chart (the reference's only multiplicity: re-installing with
 * another nameOverride, README.md:72) rendered together: all Secret payloads of the fleet
 * go through ONE GPU batch. out_off receives n_releases+1 offsets into *out (kv_free). */
int kv_fleet_template(const kv_chart *chart, int n_releases, const char *const *release_names,
                      const char *ns, const kv_overrides *overrides, char **out, int64_t *out_off);
/* `helm lint` equivalent (BASELINE.json config 1). *messages: newline-separated, kv_free. */
int kv_helm_lint(const char *chart_dir, const kv_overrides *overrides, int *ok, char **messages,
                 size_t *len);
void kv_free(void *p);

/* Select the CUDA device the host-buffer entry points use (default 0). */
int kv_set_device(int device);
/* Release the per-device staging buffers and streams held by the host-buffer path. */
void kv_shutdown(void);

/* Shape of the host-buffer pipeline (no counterpart in the reference): kv_b64enc_batch and
 * kv_b64dec_batch cut a batch into chunks of about chunk_bytes input bytes (default 16 MiB) and
 * keep `slots` of them in flight (default 4, at most 8), each slot owning device buffers of one
 * chunk.  0 leaves a setting as it is.  Takes effect from the next batch; results do not depend
 * on it, only speed and device-memory use do.                                                */
int kv_host_pipeline_config(int64_t chunk_bytes, int slots);

/* Page-locked host memory for the buffers handed to kv_b64enc_batch / kv_b64dec_batch.  (No
 * counterpart in the reference.)  The host-buffer entries accept any memory, but from ordinary
 * pageable memory every copy is staged by the driver and the call runs about six times slower
 * (bench fleet: 112 ms instead of 18 ms per step) -- slower than the CPU.  A binding that has to
 * copy its strings out of a managed heap anyway (cgo's C.CBytes, JNI's Get...Region) should copy
 * them into a buffer from here.  Returns NULL when there is no CUDA device or no memory;
 * kv_pinned_free(NULL) is a no-op.                                                          */
void *kv_pinned_alloc(size_t bytes);
void kv_pinned_free(void *p);

#ifdef __cplusplus
}
#endif
#endif /* KVEDGE_B200_H_ */
