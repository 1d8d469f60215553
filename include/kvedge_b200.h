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
 * offsets, kv_b64enc_run_dev runs the encode kernel over it.  plan + run == batch_dev.   */
int kv_b64enc_plan_dev(const int64_t *d_in_off, int64_t n_units, int64_t total_in_bytes,
                       const int64_t *d_out_off, void *d_scratch, size_t scratch_bytes,
                       void *stream);
int kv_b64enc_run_dev(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                      int64_t total_in_bytes, char *d_out, void *d_scratch,
                      size_t scratch_bytes, void *stream);

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

/* Select the CUDA device the host-buffer entry points use (default 0). */
int kv_set_device(int device);
/* Release the per-device staging buffers and streams held by the host-buffer path. */
void kv_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* KVEDGE_B200_H_ */
