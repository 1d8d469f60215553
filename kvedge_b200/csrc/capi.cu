// extern "C" boundary of libkvedge_b200.so (declared in include/kvedge_b200.h).
//
// Two families:
//   *_dev        device-resident packed batches, asynchronous on the caller's stream;
//   host-buffer  what a template engine calls (Sprig `b64enc(string) string`, reference
//                sites aziot-edge-runtime-config-secret.yaml:6 and
//                aziot-edge-vm-cloud-init-secret.yaml:6): copies in, encodes on the GPU,
//                copies out, chunked over three streams so H2D / kernel / D2H overlap.
// There is no CPU implementation behind any of them.
#include "../../include/kvedge_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "b64_kernels.cuh"

namespace {

thread_local std::string g_err;

int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

int cuda_fail(cudaError_t e, const char *what) {
    if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ||
        e == cudaErrorInitializationError)
        return fail(KV_ENODEVICE, "%s: %s (no CUDA device; this library has no CPU fallback)",
                    what, cudaGetErrorString(e));
    return fail(KV_ECUDA, "%s: %s", what, cudaGetErrorString(e));
}

#define KV_CUDA(call)                                         \
    do {                                                      \
        cudaError_t e_ = (call);                              \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call);   \
    } while (0)

int g_sm_count[64];
std::mutex g_mu;

int sm_count_of_current_device(int *out) {
    int dev = 0;
    KV_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(KV_EINVAL, "device index %d out of range", dev);
    static std::mutex mu;   // the *_dev entries may come from any thread
    std::lock_guard<std::mutex> lk(mu);
    if (g_sm_count[dev] == 0) {
        int n = 0;
        KV_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
        g_sm_count[dev] = n;
    }
    *out = g_sm_count[dev];
    return KV_OK;
}

// ------------------------------------------------------------------ host-buffer pipeline
constexpr int kSlots = 8;       // slots that exist; n_slots() of them are used
constexpr int64_t kChunkBytesDefault = 16ll << 20;  // input bytes per pipeline chunk

// Chunk size and slots in use: kv_host_pipeline_config() (an explicit call, under g_mu).  The
// diagnostic knobs KVEDGE_TRACE and KVEDGE_E2E_DIAG exist in the tools build only
// (-DKVEDGE_TOOLS, libkvedge_b200_tools.so): the product library reads no environment variable,
// so what a caller gets never depends on its environment.
int g_n_slots = 4;
int64_t g_chunk_bytes = kChunkBytesDefault;
int n_slots() { return g_n_slots; }
int64_t chunk_bytes() { return g_chunk_bytes; }
#ifdef KVEDGE_TOOLS
int env_int(const char *name, int dflt) {
    const char *e = getenv(name);
    return e && atoi(e) > 0 ? atoi(e) : dflt;
}
#endif

// A slot owns the device buffers of one chunk in flight.  Three streams carry the batch --
// uploads, kernels, downloads -- and events chain a chunk through them, so the host submits the
// whole batch without ever waiting for a download (waiting on per-slot "done" events let the
// upload engine idle until the host woke up: measured 20.1 ms against a 17.3 ms copy-only floor).
struct Slot {
    cudaEvent_t up = nullptr;     // uploads of the chunk finished (host may rewrite h_ctl)
    cudaEvent_t ran = nullptr;    // kernel finished
    cudaEvent_t down = nullptr;   // downloads finished (device buffers free again)
    bool used = false;
    char *d_in = nullptr;   size_t cap_in = 0;
    char *d_out = nullptr;  size_t cap_out = 0;
    // control block, same layout on both sides so that ONE copy ships it:
    //   in_off[n+1] | out_off[n+1] | out_len[n] | status[n] | pad to 256 | scratch (header, tile table)
    char *d_ctl = nullptr;  size_t cap_ctl = 0;
    char *h_ctl = nullptr;  size_t cap_hctl = 0;  // pinned
};

struct HostCtx {
    int device = 0;
    bool ready = false;
    cudaStream_t s_up = nullptr, s_run = nullptr, s_down = nullptr;
    Slot slot[kSlots];
};
HostCtx g_ctx;

int grow(char **p, size_t *cap, size_t need) {
    if (need <= *cap) return KV_OK;
    if (*p) KV_CUDA(cudaFree(*p));
    *p = nullptr;
    *cap = 0;
    size_t want = std::max(need, (size_t)1 << 20);
    want = (want + (want >> 2) + 255) / 256 * 256;
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, want);
    if (e != cudaSuccess) return fail(KV_ENOMEM, "cudaMalloc(%zu): %s", want, cudaGetErrorString(e));
    *p = static_cast<char *>(q);
    *cap = want;
    return KV_OK;
}

int grow_pinned(char **p, size_t *cap, size_t need) {
    if (need <= *cap) return KV_OK;
    if (*p) KV_CUDA(cudaFreeHost(*p));
    *p = nullptr;
    *cap = 0;
    size_t want = std::max(need, (size_t)1 << 16);
    want = (want + (want >> 2) + 255) / 256 * 256;
    void *q = nullptr;
    cudaError_t e = cudaMallocHost(&q, want);
    if (e != cudaSuccess) return fail(KV_ENOMEM, "cudaMallocHost(%zu): %s", want, cudaGetErrorString(e));
    *p = static_cast<char *>(q);
    *cap = want;
    return KV_OK;
}

int ctx_init() {
    if (g_ctx.ready) return KV_OK;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
    if (n <= 0) return fail(KV_ENODEVICE, "no CUDA device (this library has no CPU fallback)");
    KV_CUDA(cudaSetDevice(g_ctx.device));
    KV_CUDA(cudaStreamCreateWithFlags(&g_ctx.s_up, cudaStreamNonBlocking));
    KV_CUDA(cudaStreamCreateWithFlags(&g_ctx.s_run, cudaStreamNonBlocking));
    KV_CUDA(cudaStreamCreateWithFlags(&g_ctx.s_down, cudaStreamNonBlocking));
    for (int i = 0; i < kSlots; ++i) {
        KV_CUDA(cudaEventCreateWithFlags(&g_ctx.slot[i].up, cudaEventDisableTiming));
        KV_CUDA(cudaEventCreateWithFlags(&g_ctx.slot[i].ran, cudaEventDisableTiming));
        KV_CUDA(cudaEventCreateWithFlags(&g_ctx.slot[i].down, cudaEventDisableTiming));
    }
    g_ctx.ready = true;
    return KV_OK;
}

// KVEDGE_TRACE=1: per-chunk device timeline of a host batch on stderr (tools/ only).
struct ChunkTrace { cudaEvent_t e[4]; double host[5]; };  // before H2D, after H2D, after kernel, after D2H; host clock at each mark
std::vector<ChunkTrace> g_trace;
#ifdef KVEDGE_TOOLS
bool trace_on() {
    static const bool v = env_int("KVEDGE_TRACE", 0) > 0;
    return v;
}
#else
constexpr bool trace_on() { return false; }
#endif
void trace_mark(int which, cudaStream_t st) {
    if (!trace_on()) return;
    if (which == 0) {
        ChunkTrace t;
        for (auto &e : t.e) cudaEventCreate(&e);
        g_trace.push_back(t);
    }
    cudaEventRecord(g_trace.back().e[which], st);
    g_trace.back().host[which] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
void trace_dump() {
    if (!trace_on() || g_trace.empty()) return;
    for (size_t i = 0; i < g_trace.size(); ++i) {
        float t[4];
        for (int k = 0; k < 4; ++k) cudaEventElapsedTime(&t[k], g_trace[0].e[0], g_trace[i].e[k]);
        const double h0 = g_trace[0].host[0];
        fprintf(stderr, "chunk %2zu: h2d %7.3f..%7.3f  kernel ..%7.3f  d2h ..%7.3f ms   host submits at %7.3f %7.3f %7.3f %7.3f\n", i,
                t[0], t[1], t[2], t[3], g_trace[i].host[0] - h0, g_trace[i].host[1] - h0, g_trace[i].host[2] - h0, g_trace[i].host[3] - h0);
    }
    for (auto &c : g_trace)
        for (auto &e : c.e) cudaEventDestroy(e);
    g_trace.clear();
}

// Run units [lo, hi) of a host packed batch through one slot. mode 0 = encode, 1 = decode.
// The tile table is built here on the host (every offset is known) and travels with the
// offsets in a single copy: per chunk the GPU sees one control copy, one data copy, the
// encode (or decode) kernel and one copy back.
// lone: the chunk is the whole batch (a single render, a handful of Secrets).  Nothing can
// overlap then, so everything goes down ONE stream with no events in between, and a small input
// rides in the control block's copy: latency is all such a call has.
int run_chunk(Slot &s, int mode, const char *in, const int64_t *in_off, char *out,
              const int64_t *out_off, int64_t lo, int64_t hi, int64_t *out_len,
              int64_t *status, int sm_count, bool lone) {
    int rc;
    // the previous chunk in this slot: its uploads must be over before h_ctl is rewritten
    // (long past), and its downloads before the device buffers are written again (ordered on
    // the device, not by the host)
    if (s.used) KV_CUDA(cudaEventSynchronize(s.up));
    const int64_t n = hi - lo;
    const int64_t ibeg = in_off[lo], iend = in_off[hi];
    const int64_t obeg = out_off[lo], oend = out_off[hi];
    const int64_t in_bytes = iend - ibeg, out_bytes = oend - obeg;
    // keep each unit's alignment (mod 16) identical on the device and on the host side
    // and give the device side of each copy the host side's offset inside a 4 KiB page: the copy
    // engines run measurably below their rate when source and destination are skewed
    const bool inline_input = lone && in_bytes <= (48 << 10);
    const int64_t ishift = inline_input ? 16 : (int64_t)(reinterpret_cast<uintptr_t>(in + ibeg) & 4095u);
    const int64_t oshift = (int64_t)(reinterpret_cast<uintptr_t>(out + obeg) & 4095u);
    const int tile = mode == 0 ? kvedge::kEncTileIn : kvedge::dec_tile_in();
    const kvedge::ScratchLayout L = kvedge::scratch_layout(n, in_bytes + ishift, tile);
    const size_t off_in = 0, off_out = (size_t)(n + 1) * 8, off_len = off_out + (size_t)(n + 1) * 8;
    const size_t off_status = off_len + (size_t)n * 8;
    const size_t off_scratch = (off_status + (size_t)n * 8 + 255) / 256 * 256;
    const size_t ctl_total = off_scratch + L.total + (inline_input ? (size_t)in_bytes + 512 : 0);
    if (!inline_input && (rc = grow(&s.d_in, &s.cap_in, (size_t)(in_bytes + ishift + 32)))) return rc;
    if ((rc = grow(&s.d_out, &s.cap_out, (size_t)(out_bytes + oshift + 32)))) return rc;
    if ((rc = grow(&s.d_ctl, &s.cap_ctl, ctl_total))) return rc;
    if ((rc = grow_pinned(&s.h_ctl, &s.cap_hctl, ctl_total))) return rc;
    int64_t *h_in_off = reinterpret_cast<int64_t *>(s.h_ctl + off_in);
    int64_t *h_out_off = reinterpret_cast<int64_t *>(s.h_ctl + off_out);
    int64_t *h_len = reinterpret_cast<int64_t *>(s.h_ctl + off_len);
    int64_t *h_status = reinterpret_cast<int64_t *>(s.h_ctl + off_status);
    for (int64_t k = 0; k <= n; ++k) {
        h_in_off[k] = in_off[lo + k] - ibeg + ishift;
        h_out_off[k] = out_off[lo + k] - obeg + oshift;
    }
    const int small_max = mode == 0 ? kvedge::enc_small_max(L) : kvedge::dec_small_max();
    const int64_t n_tiles = kvedge::plan_tiles_host(h_in_off, h_out_off, n, tile, small_max, mode, L,
                                                    s.h_ctl + off_scratch,
                                                    reinterpret_cast<const int64_t *>(s.d_ctl + off_out),
                                                    h_len, h_status);
    size_t ctl_used = off_scratch + L.desc_off + (size_t)n_tiles * sizeof(kvedge::TileDesc);
    const char *d_input = s.d_in;
    if (inline_input) {
        const size_t data_base = (ctl_used + 255) / 256 * 256;
        memcpy(s.h_ctl + data_base + ishift, in + ibeg, (size_t)in_bytes);
        d_input = s.d_ctl + data_base;
        ctl_used = data_base + (size_t)ishift + (size_t)in_bytes;
    }
    cudaStream_t s_run = g_ctx.s_run;
    cudaStream_t s_up = lone ? s_run : g_ctx.s_up, s_down = lone ? s_run : g_ctx.s_down;
    if (s.used) KV_CUDA(cudaStreamWaitEvent(s_up, s.down, 0));
    trace_mark(0, s_up);
    KV_CUDA(cudaMemcpyAsync(s.d_ctl, s.h_ctl, ctl_used, cudaMemcpyHostToDevice, s_up));
    if (in_bytes > 0 && !inline_input) {
        // An upload whose host address is not 256-byte aligned runs ~10 % slower and drags the
        // concurrent download with it (tools/pcie_peak.py): send the ragged head on its own.
        size_t head = (size_t)((256u - (reinterpret_cast<uintptr_t>(in + ibeg) & 255u)) & 255u);
        if (head > (size_t)in_bytes) head = (size_t)in_bytes;
        if (head) KV_CUDA(cudaMemcpyAsync(s.d_in + ishift, in + ibeg, head, cudaMemcpyHostToDevice, s_up));
        if ((size_t)in_bytes > head)
            KV_CUDA(cudaMemcpyAsync(s.d_in + ishift + head, in + ibeg + head, (size_t)in_bytes - head, cudaMemcpyHostToDevice, s_up));
    }
    trace_mark(1, s_up);
    KV_CUDA(cudaEventRecord(s.up, s_up));
    if (!lone) KV_CUDA(cudaStreamWaitEvent(s_run, s.up, 0));
    const int64_t *d_in_off = reinterpret_cast<const int64_t *>(s.d_ctl + off_in);
    const int64_t *d_out_off = reinterpret_cast<const int64_t *>(s.d_ctl + off_out);
    int64_t *d_len = reinterpret_cast<int64_t *>(s.d_ctl + off_len);
    int64_t *d_status = reinterpret_cast<int64_t *>(s.d_ctl + off_status);
    cudaError_t e = cudaSuccess;
#ifdef KVEDGE_TOOLS
    static const int diag = env_int("KVEDGE_E2E_DIAG", 0);  // 1 = no kernel, 2 = also no upload->download ordering
#else
    constexpr int diag = 0;
#endif
    if (n_tiles > 0 && diag == 0) {
        if (mode == 0)
            e = kvedge::launch_b64enc_run(reinterpret_cast<const uint8_t *>(d_input), d_in_off, n,
                                          in_bytes + ishift, s.d_out, s.d_ctl + off_scratch, L.total,
                                          s_run, sm_count);
        else
            e = kvedge::launch_b64dec_run(d_input, d_in_off, n, in_bytes + ishift,
                                          reinterpret_cast<uint8_t *>(s.d_out), d_out_off, d_len, d_status,
                                          s.d_ctl + off_scratch, L.total, s_run, sm_count);
    }
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    trace_mark(2, s_run);
    KV_CUDA(cudaEventRecord(s.ran, s_run));
    if (diag != 2 && !lone) KV_CUDA(cudaStreamWaitEvent(s_down, s.ran, 0));
    if (out_bytes > 0)
        KV_CUDA(cudaMemcpyAsync(out + obeg, s.d_out + oshift, (size_t)out_bytes, cudaMemcpyDeviceToHost, s_down));
    if (mode == 1) {
        KV_CUDA(cudaMemcpyAsync(out_len + lo, d_len, (size_t)n * 8, cudaMemcpyDeviceToHost, s_down));
        KV_CUDA(cudaMemcpyAsync(status + lo, d_status, (size_t)n * 8, cudaMemcpyDeviceToHost, s_down));
    }
    trace_mark(3, s_down);
    KV_CUDA(cudaEventRecord(s.down, s_down));
    s.used = true;
    return KV_OK;
}

int run_host_batch(int mode, const char *in, const int64_t *in_off, int64_t n_units, char *out,
                   const int64_t *out_off, int64_t *out_len, int64_t *status) {
    if (n_units < 0) return fail(KV_EINVAL, "n_units < 0");
    if (n_units == 0) return KV_OK;
    if (!in_off || !out_off) return fail(KV_EINVAL, "null offsets");
    for (int64_t i = 0; i < n_units; ++i) {
        const int64_t len = in_off[i + 1] - in_off[i];
        if (len < 0) return fail(KV_EINVAL, "in_off is not non-decreasing at unit %lld", (long long)i);
        const int64_t need = mode == 0 ? (int64_t)kv_b64enc_len((size_t)len)
                                       : (int64_t)kv_b64dec_maxlen((size_t)len);
        if (out_off[i + 1] - out_off[i] < need)
            return fail(KV_EINVAL, "unit %lld: output capacity %lld < %lld", (long long)i,
                        (long long)(out_off[i + 1] - out_off[i]), (long long)need);
    }
    if (in_off[n_units] > in_off[0] && !in) return fail(KV_EINVAL, "null input");
    if (out_off[n_units] > out_off[0] && !out) return fail(KV_EINVAL, "null output");
    std::lock_guard<std::mutex> lk(g_mu);
    int rc = ctx_init();
    if (rc) return rc;
    KV_CUDA(cudaSetDevice(g_ctx.device));
    int sms = 0;
    if ((rc = sm_count_of_current_device(&sms))) return rc;
    int64_t lo = 0;
    int turn = 0;
    bool all_lone = true;
    while (lo < n_units) {
        int64_t hi = lo + 1;
        while (hi < n_units && in_off[hi + 1] - in_off[lo] <= chunk_bytes()) ++hi;
        all_lone = all_lone && lo == 0 && hi == n_units;
        rc = run_chunk(g_ctx.slot[turn % n_slots()], mode, in, in_off, out, out_off, lo, hi, out_len,
                       status, sms, lo == 0 && hi == n_units);
        if (rc) break;
        lo = hi;
        ++turn;
    }
    // everything a chunk does ends in the download stream (a lone chunk: in the kernel stream)
    const bool was_lone = all_lone;
    cudaError_t e2 = cudaStreamSynchronize(g_ctx.s_run);
    cudaError_t e1 = was_lone ? cudaSuccess : cudaStreamSynchronize(g_ctx.s_down);
    cudaError_t e3 = was_lone ? cudaSuccess : cudaStreamSynchronize(g_ctx.s_up);
    if (!rc && e1 != cudaSuccess) rc = cuda_fail(e1, "download stream");
    if (!rc && e2 != cudaSuccess) rc = cuda_fail(e2, "kernel stream");
    if (!rc && e3 != cudaSuccess) rc = cuda_fail(e3, "upload stream");
    trace_dump();
    return rc;
}

}  // namespace

extern "C" {

// Pinned staging for kv_fleet_template (render_capi.cpp): the fleet's packed inputs and their
// encodings, kept between calls -- fresh pageable buffers of this size cost more in page faults
// and staged copies than the encode itself.  Not part of the public header.
static char *g_stage_in = nullptr, *g_stage_out = nullptr;
static size_t g_stage_in_cap = 0, g_stage_out_cap = 0;
int kv_stage_get_(size_t in_bytes, size_t out_bytes, char **in, char **out) {
    std::lock_guard<std::mutex> lk(g_mu);
    int rc = ctx_init();
    if (rc) return rc;
    KV_CUDA(cudaSetDevice(g_ctx.device));
    if ((rc = grow_pinned(&g_stage_in, &g_stage_in_cap, in_bytes + 256))) return rc;
    if ((rc = grow_pinned(&g_stage_out, &g_stage_out_cap, out_bytes + 256))) return rc;
    *in = g_stage_in;
    *out = g_stage_out;
    return KV_OK;
}

// shared with render_capi.cpp (not part of the public header)
int kv_set_last_error_(int code, const char *msg) {
    g_err = msg ? msg : "";
    return code;
}

int kv_version(void) { return 200; }  // 0.2.0: kv_b64_batch_status_dev, kv_host_pipeline_config

const char *kv_last_error(void) { return g_err.c_str(); }

int kv_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        (void)cudaGetLastError();
        return 0;
    }
    return n;
}

size_t kv_b64enc_len(size_t n) { return (n + 2) / 3 * 4; }
size_t kv_b64dec_maxlen(size_t n) { return n / 4 * 3; }

size_t kv_b64_scratch_bytes(int64_t n_units, int64_t total_in_bytes) {
    // sized for the smaller of the two tile sizes, so one buffer serves encode and decode
    return kvedge::scratch_layout(n_units, total_in_bytes, kvedge::kMinTileIn).total;
}

int kv_b64enc_batch_dev(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                        int64_t total_in_bytes, char *d_out, const int64_t *d_out_off,
                        void *d_scratch, size_t scratch_bytes, void *stream) {
    if (n_units < 0 || total_in_bytes < 0) return fail(KV_EINVAL, "negative size");
    if (n_units == 0) return KV_OK;
    if (!d_in_off || !d_out_off || !d_scratch) return fail(KV_EINVAL, "null pointer");
    size_t need = kvedge::scratch_layout(n_units, total_in_bytes, kvedge::kEncTileIn).total;
    if (scratch_bytes < need)
        return fail(KV_ENOMEM, "scratch too small: %zu < %zu (kv_b64_scratch_bytes)", scratch_bytes, need);
    int sms = 0;
    int rc = sm_count_of_current_device(&sms);
    if (rc) return rc;
    cudaError_t e = kvedge::launch_b64enc(d_in, d_in_off, n_units, total_in_bytes, d_out, d_out_off,
                                          d_scratch, scratch_bytes, static_cast<cudaStream_t>(stream), sms);
    if (e != cudaSuccess) return cuda_fail(e, "b64enc launch");
    return KV_OK;
}

int kv_b64enc_plan_dev(const int64_t *d_in_off, int64_t n_units, int64_t total_in_bytes,
                       const int64_t *d_out_off, void *d_scratch, size_t scratch_bytes,
                       void *stream) {
    if (n_units < 0 || total_in_bytes < 0) return fail(KV_EINVAL, "negative size");
    if (n_units == 0) return KV_OK;
    if (!d_in_off || !d_out_off || !d_scratch) return fail(KV_EINVAL, "null pointer");
    size_t need = kvedge::scratch_layout(n_units, total_in_bytes, kvedge::kEncTileIn).total;
    if (scratch_bytes < need)
        return fail(KV_ENOMEM, "scratch too small: %zu < %zu (kv_b64_scratch_bytes)", scratch_bytes, need);
    cudaError_t e = kvedge::launch_b64enc_plan(d_in_off, n_units, total_in_bytes, d_out_off, d_scratch,
                                               scratch_bytes, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "b64enc plan launch");
    return KV_OK;
}

int kv_b64enc_run_dev(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                      int64_t total_in_bytes, char *d_out, void *d_scratch,
                      size_t scratch_bytes, void *stream) {
    if (n_units < 0 || total_in_bytes < 0) return fail(KV_EINVAL, "negative size");
    if (n_units == 0) return KV_OK;
    if (!d_in_off || !d_scratch) return fail(KV_EINVAL, "null pointer");
    size_t need = kvedge::scratch_layout(n_units, total_in_bytes, kvedge::kEncTileIn).total;
    if (scratch_bytes < need)
        return fail(KV_ENOMEM, "scratch too small: %zu < %zu (kv_b64_scratch_bytes)", scratch_bytes, need);
    int sms = 0;
    int rc = sm_count_of_current_device(&sms);
    if (rc) return rc;
    cudaError_t e = kvedge::launch_b64enc_run(d_in, d_in_off, n_units, total_in_bytes, d_out, d_scratch,
                                              scratch_bytes, static_cast<cudaStream_t>(stream), sms);
    if (e != cudaSuccess) return cuda_fail(e, "b64enc run launch");
    return KV_OK;
}

int kv_b64_batch_status_dev(const void *d_scratch, void *stream) {
    if (!d_scratch) return fail(KV_EINVAL, "null scratch");
    int header[4] = {0, 0, 0, 0};
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    KV_CUDA(cudaMemcpyAsync(header, d_scratch, sizeof header, cudaMemcpyDeviceToHost, st));
    KV_CUDA(cudaStreamSynchronize(st));
    if (header[3] != 0)
        return fail(KV_EINVAL, "the batch was truncated: total_in_bytes understated d_in_off[n_units], the tile table "
                               "held %d tiles and units beyond them were not processed", header[0]);
    return KV_OK;
}

int kv_b64dec_batch_dev(const char *d_in, const int64_t *d_in_off, int64_t n_units,
                        int64_t total_in_bytes, uint8_t *d_out, const int64_t *d_out_off,
                        int64_t *d_out_len, int64_t *d_status, void *d_scratch,
                        size_t scratch_bytes, void *stream) {
    if (n_units < 0 || total_in_bytes < 0) return fail(KV_EINVAL, "negative size");
    if (n_units == 0) return KV_OK;
    if (!d_in_off || !d_out_off || !d_scratch || !d_out_len || !d_status)
        return fail(KV_EINVAL, "null pointer");
    size_t need = kvedge::scratch_layout(n_units, total_in_bytes, kvedge::dec_tile_in()).total;
    if (scratch_bytes < need)
        return fail(KV_ENOMEM, "scratch too small: %zu < %zu (kv_b64_scratch_bytes)", scratch_bytes, need);
    int sms = 0;
    int rc = sm_count_of_current_device(&sms);
    if (rc) return rc;
    cudaError_t e = kvedge::launch_b64dec(d_in, d_in_off, n_units, total_in_bytes, d_out, d_out_off,
                                          d_out_len, d_status, d_scratch, scratch_bytes,
                                          static_cast<cudaStream_t>(stream), sms);
    if (e != cudaSuccess) return cuda_fail(e, "b64dec launch");
    return KV_OK;
}

int kv_b64enc_batch(const uint8_t *in, const int64_t *in_off, int64_t n_units, char *out,
                    const int64_t *out_off) {
    return run_host_batch(0, reinterpret_cast<const char *>(in), in_off, n_units, out, out_off,
                          nullptr, nullptr);
}

int kv_b64enc(const uint8_t *in, size_t n, char *out) {
    if (n == 0) return KV_OK;
    const int64_t io[2] = {0, (int64_t)n};
    const int64_t oo[2] = {0, (int64_t)kv_b64enc_len(n)};
    return kv_b64enc_batch(in, io, 1, out, oo);
}

int kv_b64dec_batch(const char *in, const int64_t *in_off, int64_t n_units, uint8_t *out,
                    const int64_t *out_off, int64_t *out_len, int64_t *status) {
    if (n_units > 0 && (!out_len || !status)) return fail(KV_EINVAL, "null out_len/status");
    return run_host_batch(1, in, in_off, n_units, reinterpret_cast<char *>(out), out_off, out_len,
                          status);
}

int kv_b64dec(const char *in, size_t n, uint8_t *out, size_t *out_n, size_t *err_off) {
    if (out_n) *out_n = 0;
    if (n == 0) return KV_OK;
    const int64_t io[2] = {0, (int64_t)n};
    const int64_t oo[2] = {0, (int64_t)kv_b64dec_maxlen(n)};
    int64_t len = 0, st = -1;
    int rc = kv_b64dec_batch(in, io, 1, out, oo, &len, &st);
    if (rc) return rc;
    if (out_n) *out_n = (size_t)len;
    if (st >= 0) {
        if (err_off) *err_off = (size_t)st;
        return fail(KV_ECORRUPT, "illegal base64 data at input byte %lld", (long long)st);
    }
    return KV_OK;
}

int kv_host_pipeline_config(int64_t chunk_bytes_, int slots) {
    if (chunk_bytes_ < 0 || slots < 0 || slots > kSlots)
        return fail(KV_EINVAL, "chunk_bytes must be >= 0 and slots within 0..%d", kSlots);
    std::lock_guard<std::mutex> lk(g_mu);
    if (chunk_bytes_ > 0) g_chunk_bytes = std::max<int64_t>(chunk_bytes_, 64 << 10);
    if (slots > 0) g_n_slots = std::max(slots, 2);
    return KV_OK;
}

int kv_set_device(int device) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_ctx.ready && device != g_ctx.device)
        return fail(KV_EINVAL, "kv_set_device after first use; call kv_shutdown() first");
    g_ctx.device = device;
    return KV_OK;
}

void *kv_pinned_alloc(size_t bytes) {
    void *p = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (ctx_init() != KV_OK) return nullptr;
        if (cudaSetDevice(g_ctx.device) != cudaSuccess) return nullptr;
    }
    cudaError_t e = cudaMallocHost(&p, bytes ? bytes : 1);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        fail(KV_ENOMEM, "cudaMallocHost(%zu): %s", bytes, cudaGetErrorString(e));
        return nullptr;
    }
    return p;
}

void kv_pinned_free(void *p) {
    if (p) (void)cudaFreeHost(p);
}

void kv_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_ctx.ready) return;
    cudaSetDevice(g_ctx.device);
    cudaDeviceSynchronize();
    for (int i = 0; i < kSlots; ++i) {
        Slot &s = g_ctx.slot[i];
        if (s.d_in) cudaFree(s.d_in);
        if (s.d_out) cudaFree(s.d_out);
        if (s.d_ctl) cudaFree(s.d_ctl);
        if (s.h_ctl) cudaFreeHost(s.h_ctl);
        if (s.up) cudaEventDestroy(s.up);
        if (s.ran) cudaEventDestroy(s.ran);
        if (s.down) cudaEventDestroy(s.down);
        s = Slot();
    }
    if (g_stage_in) cudaFreeHost(g_stage_in);
    if (g_stage_out) cudaFreeHost(g_stage_out);
    g_stage_in = g_stage_out = nullptr;
    g_stage_in_cap = g_stage_out_cap = 0;
    if (g_ctx.s_up) cudaStreamDestroy(g_ctx.s_up);
    if (g_ctx.s_run) cudaStreamDestroy(g_ctx.s_run);
    if (g_ctx.s_down) cudaStreamDestroy(g_ctx.s_down);
    g_ctx.s_up = g_ctx.s_run = g_ctx.s_down = nullptr;
    g_ctx.ready = false;
    kvedge::reset_launch_caches();
}

}  // extern "C"
