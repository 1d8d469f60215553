// Internal launcher interface of the base64 kernels (C++ linkage; the C ABI lives in capi.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace kvedge {

// Input bytes per encode tile: 4 rows of 256 threads x 12 bytes. A multiple of 12, so
// every tile of a unit starts on a 3-byte group and its output on a 16-byte step.
constexpr int kEncTileIn = 12288;
constexpr int kEncTileOut = kEncTileIn / 3 * 4;
// Input characters per decode tile (one ring stage of the streamed decode kernel).
// The default; dec_tile_in() is what the planners and the kernel actually use (a diagnostic
// variant may pick another size, never below kMinTileIn).
constexpr int kDecTileIn = 16384;
constexpr int kMinTileIn = 8192;  // sizes the scratch
int dec_tile_in();
int dec_small_max();   // units up to this many characters are grouped into shared tiles (0: never)
constexpr int kThreads = 256;
constexpr int kPlanBlock = 1024;  // units per block in the plan kernels
constexpr int kMaxGroupUnits = 64;  // most units one group tile may hold (see plan_count)

// One tile of work: a contiguous piece of one unit.
struct __align__(32) TileDesc {
    int64_t in_pos;   // absolute byte offset in the input blob
    int64_t out_pos;  // absolute byte offset in the output blob
    int32_t len;      // input bytes in this tile (<= tile size)
    int32_t unit;     // unit index
    int32_t flags;    // bit0: last tile of its unit; bit1: group tile (see plan_count)
    int32_t pad;      // group tile: number of consecutive units, starting at `unit`, it holds
};

// Scratch layout (device memory owned by the caller of the *_dev entry points):
//   [0, 256)                 header: int32 n_tiles
//   [256, 256 + 4*nb)        per-plan-block tile counts (nb = ceil(n_units / kPlanBlock))
//   [desc_off, ...)          TileDesc[max_tiles]
struct ScratchLayout {
    size_t block_sums_off;
    size_t desc_off;
    size_t total;
    int64_t max_tiles;
    int64_t n_plan_blocks;
};
ScratchLayout scratch_layout(int64_t n_units, int64_t total_in_bytes, int tile_in);

cudaError_t launch_b64enc(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                          int64_t total_in_bytes, char *d_out, const int64_t *d_out_off,
                          void *d_scratch, size_t scratch_bytes, cudaStream_t stream, int sm_count);

// The two halves of launch_b64enc: build the tile table once, run it many times (same
// offsets, new bytes).
cudaError_t launch_b64enc_plan(const int64_t *d_in_off, int64_t n_units, int64_t total_in_bytes,
                               const int64_t *d_out_off, void *d_scratch, size_t scratch_bytes,
                               cudaStream_t stream);
cudaError_t launch_b64enc_run(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                              int64_t total_in_bytes, char *d_out, void *d_scratch,
                              size_t scratch_bytes, cudaStream_t stream, int sm_count);

cudaError_t launch_b64dec(const char *d_in, const int64_t *d_in_off, int64_t n_units,
                          int64_t total_in_bytes, uint8_t *d_out, const int64_t *d_out_off,
                          int64_t *d_out_len, int64_t *d_status, void *d_scratch,
                          size_t scratch_bytes, cudaStream_t stream, int sm_count);

cudaError_t launch_b64dec_run(const char *d_in, const int64_t *d_in_off, int64_t n_units,
                              int64_t total_in_bytes, uint8_t *d_out, const int64_t *d_out_off,
                              int64_t *d_out_len, int64_t *d_status, void *d_scratch,
                              size_t scratch_bytes, cudaStream_t stream, int sm_count);

// Host-side twin of the plan kernels: writes header + TileDesc table into `scratch_host`
// (a host mirror of the scratch, laid out by L) and, for decode (mode 1), the initial
// out_len / status values.  Returns the number of tiles.
// small_max: enc_small_max(L) / dec_small_max().  d_out_off: where out_off will live on the
// device (group tiles make the kernels read it).
int64_t plan_tiles_host(const int64_t *in_off, const int64_t *out_off, int64_t n_units, int tile_in,
                        int small_max, int mode, const ScratchLayout &L, void *scratch_host,
                        const int64_t *d_out_off, int64_t *out_len, int64_t *status);
int enc_small_max(const ScratchLayout &L);

// Shared-memory opt-in and occupancy of every encode kernel the product launches, for the
// current device, now (otherwise the first launch of each kernel does it).
cudaError_t warm_b64enc_kernels();
// Forget the per-device launch set-up (kv_shutdown).
void reset_launch_caches();

}  // namespace kvedge
