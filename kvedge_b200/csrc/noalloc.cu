// kv_b64enc_batch_dev_static: the device-resident batched encode for callers that are not
// able to hand the library any scratch memory of their own.
//
// Who that is: code running inside a process whose device memory has already been claimed
// in full by something else (a serving engine or a framework's caching pool that sizes
// itself to "all free HBM" at start-up, so a later cudaMalloc simply fails); code on a
// latency-critical path where cudaMalloc/cudaFree are off limits because they can
// synchronise the whole device; and plugins that are only handed input and output
// pointers by their host and have no allocator to ask.  For them the tile table that the
// encode kernels need lives in a fixed array inside this library's device image.
//
// The price: the array is a fixed KV_STATIC_SCRATCH_BYTES, every GPU that loads the library
// carries it until the process exits, batches whose tile table would not fit are refused,
// and because every call uses the same array, two calls must not be in flight on one GPU
// at the same time.
//
// "Allocates nothing" is a statement about this library's calls: no cudaMalloc, cudaFree or
// allocating library routine, ever.  The CUDA driver, however, loads a module's code and
// global arrays into device memory when the module is first touched, and for this file that
// moment is the first call on a device (the cudaGetSymbolAddress below): that call needs
// room for the 8 MiB array and the kernels' code, so it belongs in start-up code, BEFORE the
// other party claims the rest of the memory and outside any stream capture.  The same first
// call does the kernels' one-time launch set-up for both the large-batch and the tiny-batch
// kernel, so later calls do no set-up of any kind.
#include "../../include/kvedge_b200.h"

#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "b64_kernels.cuh"

extern "C" int kv_set_last_error_(int code, const char *msg);  // capi.cu

namespace {

// The scratch itself.  In host code this name is only a handle for the runtime (it is not a
// device address and must never be passed to a kernel); the real address is looked up once
// per device with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_static_scratch[KV_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
void *g_scratch_addr[kMaxDevices];  // per device: device address of g_static_scratch
int g_sm_count[kMaxDevices];
std::mutex g_setup_mu;

int fail(int code, const std::string &msg) { return kv_set_last_error_(code, msg.c_str()); }

// One-time set-up for the device that is current on the calling thread.
int setup_current_device(int *dev_out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return fail(e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? KV_ENODEVICE : KV_ECUDA,
                    std::string("cudaGetDevice: ") + cudaGetErrorString(e) + " (no CPU fallback)");
    if (dev < 0 || dev >= kMaxDevices) return fail(KV_EINVAL, "device index out of range");
    std::lock_guard<std::mutex> lk(g_setup_mu);
    if (!g_scratch_addr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);  // with `dev` current: that device's copy
        if (e != cudaSuccess || !p)
            return fail(e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? KV_ENODEVICE : KV_ECUDA,
                        std::string("cudaGetSymbolAddress(static scratch): ") + cudaGetErrorString(e));
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return fail(KV_ECUDA, std::string("cudaDeviceGetAttribute: ") + cudaGetErrorString(e));
        // every launch-time set-up the encode kernels would otherwise do on first use -- the large-
        // batch and the tiny-batch kernel alike -- so that no later call has any left to do
        e = kvedge::warm_b64enc_kernels();
        if (e != cudaSuccess) return fail(KV_ECUDA, std::string("kernel set-up: ") + cudaGetErrorString(e));
        g_sm_count[dev] = sms;
        g_scratch_addr[dev] = p;
    }
    *dev_out = dev;
    return KV_OK;
}

}  // namespace

extern "C" {

int kv_b64enc_batch_dev_static(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                               int64_t total_in_bytes, char *d_out, const int64_t *d_out_off,
                               void *stream) {
    if (n_units < 0 || total_in_bytes < 0) return fail(KV_EINVAL, "negative size");
    if (n_units == 0) return KV_OK;
    if (!d_in_off || !d_out_off) return fail(KV_EINVAL, "null pointer");
    const size_t need = kvedge::scratch_layout(n_units, total_in_bytes, kvedge::kEncTileIn).total;
    if (need > (size_t)KV_STATIC_SCRATCH_BYTES)
        return fail(KV_ENOMEM,
                    "batch too large for the built-in scratch: its tile table needs " + std::to_string(need) +
                        " bytes, the static array holds " + std::to_string((size_t)KV_STATIC_SCRATCH_BYTES) +
                        " (about 32 bytes per unit plus 32 bytes per 12 KiB of input); split the batch or use "
                        "kv_b64enc_batch_dev with caller-owned scratch");
    int dev = 0;
    int rc = setup_current_device(&dev);
    if (rc) return rc;
    // Nothing of the previous call is reused: the plan kernels rewrite the header (tile count,
    // tile counter, finished-block counter), every per-block tile count and every descriptor
    // they or the encode kernel will read.
    cudaError_t e = kvedge::launch_b64enc(d_in, d_in_off, n_units, total_in_bytes, d_out, d_out_off,
                                          g_scratch_addr[dev], (size_t)KV_STATIC_SCRATCH_BYTES,
                                          static_cast<cudaStream_t>(stream), g_sm_count[dev]);
    if (e != cudaSuccess) return fail(KV_ECUDA, std::string("b64enc launch: ") + cudaGetErrorString(e));
    return KV_OK;
}

}  // extern "C"
