// Batched standard-base64 encode / decode for sm_100a (B200).
//
// The hot path of levi106/kvedge's only byte transform: Sprig `b64enc` at
//   deployment/helm/templates/aziot-edge-runtime-config-secret.yaml:6
//   deployment/helm/templates/aziot-edge-vm-cloud-init-secret.yaml:6
// and its inverse (Kubernetes decoding Secret.data.userdata for the disk read at
// deployment/helm/templates/_helper.tpl:64,73).  Semantics are Go's
// encoding/base64.StdEncoding (RFC 4648 section 4 alphabet, '=' padding, no wrapping).
//
// Shape of the work: many independent units (one Secret payload each) of ragged
// length in one packed blob.  It is HBM-bound byte work -- 1 byte read and 4/3 bytes
// written per input byte -- so the design goals are coalesced 16-byte stores, enough
// bytes in flight per SM (bulk async copies into shared-memory rings), and a sextet->ASCII
// mapping balanced over the ALU, FMA and load/store pipes.  No tensor cores: nothing here
// is a GEMM.
//
//   plan_count / plan_emit   ragged units -> tiles (TileDesc table in scratch)
//   b64enc_groups            encode: one block per SM, five producer/consumer pipelines
//   b64enc_stream            encode: one pipeline per block (tiny batches; earlier product)
//   b64dec_stream            decode of canonical input; anything else is flagged ...
//   b64dec_slow              ... and re-decoded sequentially with Go's exact semantics
#include "b64_kernels.cuh"

#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>

namespace kvedge {

// ----------------------------------------------------------------------------- layout
ScratchLayout scratch_layout(int64_t n_units, int64_t total_in_bytes, int tile_in) {
    ScratchLayout L;
    if (n_units < 0) n_units = 0;
    if (total_in_bytes < 0) total_in_bytes = 0;
    L.n_plan_blocks = (n_units + kPlanBlock - 1) / kPlanBlock;
    L.max_tiles = total_in_bytes / tile_in + n_units;
    L.block_sums_off = 256;
    size_t bs = (size_t)L.n_plan_blocks * sizeof(int32_t);
    L.desc_off = (L.block_sums_off + bs + 255) / 256 * 256;
    L.total = L.desc_off + (size_t)L.max_tiles * sizeof(TileDesc) + 256;
    return L;
}

// ----------------------------------------------------------------------------- programmatic launch
// plan_count -> plan_emit -> encode/decode kernel are launched back to back in one stream.  With
// the programmatic-serialization launch attribute a kernel may start while its predecessor is
// still running: pdl_launch_dependents() in the predecessor says "the next kernel may be
// scheduled now", pdl_wait() in the successor blocks until the predecessor has completed and its
// writes are visible.  Everything a kernel does before pdl_wait() (table fills, barrier
// initialisation) therefore overlaps the predecessor, and the launch latency disappears.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
static cudaError_t launch_after_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                       Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// ----------------------------------------------------------------------------- plan
__device__ __forceinline__ int warp_incl_scan(int v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) >= d) v += t;
    }
    return v;
}

// Exclusive scan of one int per thread across a 1024-thread block; returns the
// exclusive prefix and stores the block total in *total.
static_assert(kPlanBlock == 1024, "the block scans below assume exactly 32 warps per plan block");
__device__ __forceinline__ int block_excl_scan(int v, int *total) {
    __shared__ int warp_sums[32];
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int incl = warp_incl_scan(v);
    if (lane == 31) warp_sums[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int s = warp_sums[lane];
        int si = warp_incl_scan(s);
        warp_sums[lane] = si - s;  // exclusive prefix of each warp
        if (lane == 31) *total = si;
    }
    __syncthreads();
    return incl - v + warp_sums[wid];
}

// Tiles come in two kinds.  A unit longer than small_max is cut into pieces of tile_in bytes,
// one tile each ("solo" tiles).  Shorter units are not worth a pipeline stage of their own (a
// stage costs the same bookkeeping whether it holds 300 bytes or 12 KiB), so a run of consecutive
// short units whose first bytes fall into the same window of (tile_in - small_max) input bytes
// becomes ONE "group" tile: its span is at most tile_in bytes, and the consumer warps share its
// units.  A run also ends at every kMaxGroupUnits-th unit (hence at every plan-block boundary):
// that keeps the rule local and bounds the offsets a kernel stages per tile.
// small_max == 0 turns grouping off (the older kernels only understand solo tiles).
__host__ __device__ inline bool group_starts_at(const int64_t *in_off, int64_t u, int small_max, int window) {
    static_assert(kPlanBlock % kMaxGroupUnits == 0, "a plan block is a whole number of unit runs");
    if (u % kMaxGroupUnits == 0) return true;   // also every plan-block boundary
    if (in_off[u] - in_off[u - 1] > small_max) return true;   // the unit before is a long one
    return in_off[u] / window != in_off[u - 1] / window;
}
__host__ __device__ inline int unit_tile_count(const int64_t *in_off, int64_t u, int tile_in, int small_max) {
    const int64_t len = in_off[u + 1] - in_off[u];
    if (len <= small_max) return small_max > 0 && group_starts_at(in_off, u, small_max, tile_in - small_max) ? 1 : 0;
    return (int)((len + tile_in - 1) / tile_in);
}

__global__ void __launch_bounds__(kPlanBlock)
plan_count(const int64_t *__restrict__ in_off, int64_t n_units, int tile_in, int small_max,
           int *__restrict__ block_sums) {
    __shared__ int total;
    pdl_launch_dependents();   // plan_emit may be scheduled; it waits for our block_sums itself
    int64_t u = (int64_t)blockIdx.x * kPlanBlock + threadIdx.x;
    int cnt = 0;
    if (u < n_units) cnt = unit_tile_count(in_off, u, tile_in, small_max);
    (void)block_excl_scan(cnt, &total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// For every thread: the smallest value held by a thread with a HIGHER index in the block
// (kPlanBlock for the last one).  v is the thread's own index if it is flagged, else kPlanBlock.
__device__ __forceinline__ int block_next_flagged(int v) {
    __shared__ int warp_min[32], warp_suffix[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {   // inclusive suffix minimum inside the warp
        int t = __shfl_down_sync(0xffffffffu, v, d);
        if (lane + d < 32) v = min(v, t);
    }
    if (lane == 0) warp_min[wid] = v;
    __syncthreads();
    if (wid == 0) {
        int x = warp_min[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_down_sync(0xffffffffu, x, d);
            if (lane + d < 32) x = min(x, t);
        }
        int y = __shfl_down_sync(0xffffffffu, x, 1);
        warp_suffix[lane] = lane == 31 ? kPlanBlock : y;   // exclusive: warps after this one
    }
    __syncthreads();
    int nxt = __shfl_down_sync(0xffffffffu, v, 1);
    if (lane == 31) nxt = kPlanBlock;
    return min(nxt, warp_suffix[wid]);
}

// MODE 0: encode (out step 4 bytes per 3 in). MODE 1: decode (3 per 4) + status init.
template <int MODE>
__global__ void __launch_bounds__(kPlanBlock)
plan_emit(const int64_t *__restrict__ in_off, const int64_t *__restrict__ out_off,
          int64_t n_units, int tile_in, int small_max, const int *__restrict__ block_sums,
          int *__restrict__ header, TileDesc *__restrict__ desc, int max_tiles,
          int64_t *__restrict__ out_len, int64_t *__restrict__ status) {
    __shared__ int total;
    __shared__ int base_sh;
    __shared__ int red[32];
    pdl_launch_dependents();   // the encode/decode kernel may start its prologue
    pdl_wait();                // plan_count's block_sums
    // tiles emitted by all earlier plan blocks
    int part = 0;
    for (int b = threadIdx.x; b < (int)blockIdx.x; b += kPlanBlock) part += block_sums[b];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
    __syncthreads();
    if (threadIdx.x < 32) {
        int s = red[threadIdx.x];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
        if (threadIdx.x == 0) base_sh = s;
    }
    __syncthreads();

    int64_t u = (int64_t)blockIdx.x * kPlanBlock + threadIdx.x;
    int cnt = 0;
    int64_t ibeg = 0, len = 0, obeg = 0;
    if (u < n_units) {
        ibeg = in_off[u];
        len = in_off[u + 1] - ibeg;
        if (len < 0) len = 0;
        obeg = out_off[u];
        cnt = unit_tile_count(in_off, u, tile_in, small_max);
        if (MODE == 1 && blockIdx.y == 0) {
            out_len[u] = len / 4 * 3;
            status[u] = (len & 3) ? -2 : -1;  // -2: not canonical -> b64dec_slow decides
        }
    }
    int first = base_sh + block_excl_scan(cnt, &total);
    if (blockIdx.x == gridDim.x - 1 && blockIdx.y == 0 && threadIdx.x == 0) {
        // A caller that understated total_in_bytes sized the table too small: never write past
        // it (the batch then comes out truncated, and header[3] says so, but memory stays intact).
        header[0] = base_sh + total <= max_tiles ? base_sh + total : max_tiles;
        header[1] = 0;  // tile counter and finished-block counter of the consuming kernel
        header[2] = 0;
        header[3] = base_sh + total > max_tiles ? 1 : 0;
        // group tiles carry no output position: the kernels look each unit's up here
        const unsigned long long op = (unsigned long long)reinterpret_cast<uintptr_t>(out_off);
        header[4] = (int)(unsigned)(op & 0xffffffffull);
        header[5] = (int)(unsigned)(op >> 32);
    }
    // a group runs from its first unit to the next flagged one (a long unit or another group's start)
    const bool is_small = u < n_units && len <= small_max;
    const bool flagged = u < n_units ? (!is_small || cnt > 0) : true;
    const int next_flagged = block_next_flagged(flagged ? (int)threadIdx.x : kPlanBlock);
    if (is_small && cnt > 0 && blockIdx.y == 0) {
        int64_t last = (int64_t)blockIdx.x * kPlanBlock + next_flagged;
        if (last > n_units) last = n_units;
        TileDesc d;
        d.in_pos = ibeg;
        d.unit = (int32_t)u;
        if (last - u == 1) {
            // alone between two long units: an ordinary tile, shared by all consumer warps
            d.out_pos = obeg;
            d.len = (int32_t)len;
            d.flags = 1;
            d.pad = 0;
        } else {
            d.out_pos = 0;
            d.len = (int32_t)(in_off[last] - ibeg);
            d.flags = 2;                  // bit1: group tile; pad = number of units in it
            d.pad = (int32_t)(last - u);
        }
        if (first < max_tiles) desc[first] = d;
    }
    if (is_small) cnt = 0;   // nothing more to emit for it below
    auto emit = [&](int64_t e_ibeg, int64_t e_obeg, int64_t e_len, int64_t e_u, int e_first, int e_cnt, int k) {
        int64_t s = (int64_t)k * tile_in;
        TileDesc d;
        d.in_pos = e_ibeg + s;
        d.out_pos = e_obeg + (MODE == 0 ? s / 3 * 4 : s / 4 * 3);
        int64_t rem = e_len - s;
        d.len = (int32_t)(rem < tile_in ? rem : tile_in);
        d.unit = (int32_t)e_u;
        d.flags = (k == e_cnt - 1) ? 1 : 0;
        d.pad = 0;
        if (e_first + k < max_tiles) desc[e_first + k] = d;
    };
    // A unit of one or two tiles is written by its own thread; a long unit (a 1 MiB Secret is
    // 86 tiles) by its whole warp, or the plan would take longer than the encode.
    // When the batch is a few very long units, gridDim.y blocks share each plan block's units
    // (they all redo the scan, which is cheap) and split the long units' tiles between them.
    constexpr int kOwn = 2;
    if (cnt <= kOwn && blockIdx.y == 0)
        for (int k = 0; k < cnt; ++k) emit(ibeg, obeg, len, u, first, cnt, k);
    unsigned long_mask = __ballot_sync(0xffffffffu, cnt > kOwn);
    const int lane = threadIdx.x & 31;
    while (long_mask) {
        const int src = __ffs(long_mask) - 1;
        long_mask &= long_mask - 1;
        const int64_t b_ibeg = __shfl_sync(0xffffffffu, ibeg, src), b_obeg = __shfl_sync(0xffffffffu, obeg, src);
        const int64_t b_len = __shfl_sync(0xffffffffu, len, src), b_u = __shfl_sync(0xffffffffu, u, src);
        const int b_first = __shfl_sync(0xffffffffu, first, src), b_cnt = __shfl_sync(0xffffffffu, cnt, src);
        for (int k = lane + 32 * (int)blockIdx.y; k < b_cnt; k += 32 * (int)gridDim.y)
            emit(b_ibeg, b_obeg, b_len, b_u, b_first, b_cnt, k);
    }
}

// ----------------------------------------------------------------------------- encode
// One 3-byte group -> four RFC 4648 alphabet characters, branch and table free.
//
// np = b0<<24 | b1<<16 | b2<<8 | (don't care).  The four 6-bit values v0..v3 are moved
// into byte lanes 0..3 (first character in the lowest lane) by three masks and four
// multiplies (shifts done on the FMA pipe), already biased by -25 per lane:
//     xb = x - 0x19191919,  x = v0 | v1<<8 | v2<<16 | v3<<24.
// Per lane the character is  c = v - 25 + 6[v>25] + 75[v<=51] + 15[v<=61] + 3[v>62].
// Each comparison is a carry into bit 7 of its lane (no lane ever overflows into the
// next), and the weighted flags are folded in with multiply-high, again on the FMA
// pipe, which leaves the ALU pipe -- the scarce resource of this kernel -- with the
// byte permute, three masks and the four add+mask pairs.  Lane sums end in [43, 122],
// so plain 32-bit arithmetic never lets one lane disturb its neighbour.
// (Written in PTX so that the compiler front end cannot re-associate the sums back
// into ALU-pipe shifts and adds.)
__device__ __forceinline__ uint32_t encode_group(uint32_t np) {
    uint32_t r;
    asm("{\n\t"
        ".reg .u32 m1, m2, m3, xb, t, p1, p2, p3, p4;\n\t"
        "and.b32 m1, %1, 0x03F00000;\n\t"
        "and.b32 m2, %1, 0x000FC000;\n\t"
        "and.b32 m3, %1, 0x00003F00;\n\t"
        "shr.u32 xb, %1, 26;\n\t"
        "add.u32 xb, xb, 0xE6E6E6E7;\n\t"          // v0 - 0x19191919 (bias every lane by -25)
        "mul.hi.u32 t, m1, 0x00100000;\n\t"        // v1 -> lane 1
        "add.u32 xb, xb, t;\n\t"
        "mad.lo.u32 xb, m2, 4, xb;\n\t"            // v2 -> lane 2
        "mad.lo.u32 xb, m3, 65536, xb;\n\t"        // v3 -> lane 3
        "add.u32 p1, xb, 0x7F7F7F7F;\n\t"          // bit 7 of each lane: v > 25
        "sub.u32 p2, 0x9A9A9A9A, xb;\n\t"          //                     v <= 51
        "add.u32 p4, xb, 0x5B5B5B5B;\n\t"          // x + 0x42..: bit 7 = v > 61, bit 0 = v & 1
        "shl.b32 t, p4, 7;\n\t"
        "not.b32 p3, p4;\n\t"
        "and.b32 p4, p4, t;\n\t"                   // v > 61 and odd  <=>  v == 63
        "and.b32 p1, p1, 0x80808080;\n\t"
        "and.b32 p2, p2, 0x80808080;\n\t"
        "and.b32 p3, p3, 0x80808080;\n\t"          // v <= 61
        "and.b32 p4, p4, 0x80808080;\n\t"
        "mul.hi.u32 t, p1, 0x0C000000;\n\t"        // + 6 per flagged lane
        "add.u32 xb, xb, t;\n\t"
        "mul.hi.u32 t, p2, 0x96000000;\n\t"        // + 75
        "add.u32 xb, xb, t;\n\t"
        "mul.hi.u32 t, p3, 0x1E000000;\n\t"        // + 15
        "add.u32 xb, xb, t;\n\t"
        "mul.hi.u32 t, p4, 0x06000000;\n\t"        // + 3
        "add.u32 %0, xb, t;\n\t"
        "}"
        : "=r"(r)
        : "r"(np));
    return r;
}

// The same group through a 4096-entry shared-memory table of character pairs (12 input
// bits -> 2 characters).  Costs two LDS and six ALU operations instead of thirteen ALU
// operations: the kernel is ALU-pipe bound while the load/store pipe idles, so moving
// part of the work there shortens the critical pipe.
__device__ __forceinline__ uint32_t encode_group_lut(uint32_t np, const uint16_t *lut) {
    const uint32_t ihi = (np >> 19) & 0x1FFEu;  // byte offset of pair (v0, v1)
    const uint32_t ilo = (np >> 7) & 0x1FFEu;   // byte offset of pair (v2, v3)
    const uint32_t c01 = *reinterpret_cast<const uint16_t *>(reinterpret_cast<const char *>(lut) + ihi);
    const uint32_t c23 = *reinterpret_cast<const uint16_t *>(reinterpret_cast<const char *>(lut) + ilo);
    return __byte_perm(c01, c23, 0x5410);
}

// Fill the pair table: entry (v0<<6 | v1) = char(v0) | char(v1) << 8.
__device__ __forceinline__ void fill_pair_lut(uint16_t *lut) {
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) {
        // both sextets through the arithmetic mapping (lanes 0 and 1 of one group)
        const uint32_t np = ((uint32_t)i) << 20;  // v0 = i>>6, v1 = i&63, v2 = v3 = 0
        lut[i] = (uint16_t)(encode_group(np) & 0xFFFFu);
    }
}

// 12 input bytes in three little-endian words -> 16 output characters.  NLUT of the four
// groups go through the pair table, the others through the arithmetic mapping.
// Four 3-byte groups (each as b0<<24 | b1<<16 | b2<<8 | x) -> 16 characters.
template <int NLUT>
__device__ __forceinline__ uint4 encode_groups(uint32_t g0, uint32_t g1, uint32_t g2, uint32_t g3,
                                               const uint16_t *lut) {
    uint4 r;
#ifdef KVEDGE_TOOLS
    if (NLUT < 0) {  // diagnostic skeleton: memory traffic only, wrong bytes
        r.x = g0; r.y = g1; r.z = g2; r.w = g0 ^ g3;
        return r;
    }
#else
    static_assert(NLUT >= 0, "the traffic-only skeleton exists in the tools build only");
#endif
    r.x = NLUT >= 4 ? encode_group_lut(g0, lut) : encode_group(g0);
    r.y = NLUT >= 1 ? encode_group_lut(g1, lut) : encode_group(g1);
    r.z = NLUT >= 3 ? encode_group_lut(g2, lut) : encode_group(g2);
    r.w = NLUT >= 2 ? encode_group_lut(g3, lut) : encode_group(g3);
    return r;
}

template <int NLUT>
__device__ __forceinline__ uint4 encode12_mixed(uint32_t a0, uint32_t a1, uint32_t a2, const uint16_t *lut) {
    uint4 r;
#ifdef KVEDGE_TOOLS
    if (NLUT < 0) {  // diagnostic skeleton (tools/sweep_enc.py): memory traffic only, wrong bytes
        r.x = a0; r.y = a1; r.z = a2; r.w = a0 ^ a2;
        return r;
    }
#endif
    const uint32_t g0 = __byte_perm(a0, 0u, 0x0124), g1 = __byte_perm(a0, a1, 0x3455);
    const uint32_t g2 = __byte_perm(a1, a2, 0x2344), g3 = __byte_perm(a2, 0u, 0x1233);
    return encode_groups<NLUT>(g0, g1, g2, g3, lut);
}

__device__ __forceinline__ uint4 encode12(uint32_t a0, uint32_t a1, uint32_t a2) {
    uint4 r;
    r.x = encode_group(__byte_perm(a0, 0u, 0x0124));
    r.y = encode_group(__byte_perm(a0, a1, 0x3455));
    r.z = encode_group(__byte_perm(a1, a2, 0x2344));
    r.w = encode_group(__byte_perm(a2, 0u, 0x1233));
    return r;
}

__device__ __forceinline__ void store_chars(char *dst, uint4 v, int nchars) {
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int g = 0; g < 4; ++g) {
#pragma unroll
        for (int b = 0; b < 4; ++b)
            if (g * 4 + b < nchars) dst[g * 4 + b] = (char)((w[g] >> (8 * b)) & 0xFF);
    }
}

// Byte-wise path for the rare tiles whose word loads could leave the blob (its very
// first and last bytes) or whose destination is not even 4-byte aligned.
__device__ __noinline__ void encode_bytes(const uint8_t *p, char *q, int nvalid) {
    uint32_t a[3] = {0u, 0u, 0u};
#pragma unroll
    for (int i = 0; i < 12; ++i)
        if (i < nvalid) a[i >> 2] |= (uint32_t)p[i] << (8 * (i & 3));
    uint4 v = encode12(a[0], a[1], a[2]);
    const int rem = nvalid % 3, gi = nvalid / 3;
    if (rem) {
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int g = 0; g < 4; ++g)
            if (g == gi)
                w[g] = (rem == 1) ? ((w[g] & 0x0000FFFFu) | 0x3D3D0000u)
                                  : ((w[g] & 0x00FFFFFFu) | 0x3D000000u);
        v = make_uint4(w[0], w[1], w[2], w[3]);
    }
    store_chars(q, v, (nvalid + 2) / 3 * 4);
}

// ----------------------------------------------------------------------------- encode, streamed
// Warp-specialised variant of the encode kernel.  One producer warp per block walks the
// tile table and streams whole tiles (<= 12 KiB of one unit) into a shared-memory ring with
// one 1-D bulk async copy each (cp.async.bulk global -> shared, completion on an mbarrier);
// eight consumer warps share every stage -- warp w takes iterations w, w+8, ... of the tile,
// 384 input bytes each -- encode from shared memory and store 16 characters per lane
// straight to global memory.  The producer owns all scheduling (tile counter, descriptor
// loads, alignment), so the consumers never wait on a dependent global load, and the bytes
// in flight per SM are set by the ring depth instead of by the number of resident warps.
#ifndef KVEDGE_GROUP_NLUT
#define KVEDGE_GROUP_NLUT 3
#endif
namespace stream {

constexpr int kStageIn = kEncTileIn;       // one tile per stage
constexpr int kStageStride = (kStageIn + 32 + 15) / 16 * 16;  // 16 bytes of slack in front and behind
// b64enc_groups: behind a stage's bytes sit the offsets of the units of a group tile -- in_off[u0 ..
// u0+cnt] and out_off[u0 .. u0+cnt), each with room for one entry of alignment slack in front --
// fetched by the producer with the tile, so that no consumer ever waits on a global load for them.
constexpr int kOffsPart = (kMaxGroupUnits + 2) * 8;            // 528 bytes, a multiple of 16
constexpr int kGroupStageStride = kStageStride + 2 * kOffsPart;
constexpr int kGrabMax = 8;                // tiles taken per atomic (guided: shrinks near the end)

struct StageMeta {
    char *dst;           // where this stage's characters go
    const uint8_t *src;  // only for byte-wise stages (see flags)
    int len;             // input bytes in this stage; < 0: no more work
    uint32_t sh;         // offset of the first input byte inside the stage buffer (0..15)
    int flags;           // bit0: destination 16-byte aligned; bit1: byte-wise stage (no bulk copy)
    int pad;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// Wait with a deadline: a protocol bug must end in a trap (reported as a launch failure), never
// in a hung GPU.  (Pointer form, used by the one-pipeline kernel for tiny batches: sleeps between
// polls.)
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    unsigned ns = 32;
    while (!mbar_try_wait(bar, parity)) {
        __nanosleep(ns);
        if (ns < 256) ns *= 2;
        if (clock64() - t0 > 8000000000ll) __trap();
    }
}
// The same on 32-bit shared addresses computed once per kernel (the pointer forms make the
// compiler rebuild the shared-window address around every call).
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_a(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// try_wait with a suspend-time hint: the hardware parks the warp until the phase completes (it
// wakes at once) or the hint (in ns) runs out, so a long wait costs a handful of instructions
// instead of a poll every hundred nanoseconds.
__device__ __forceinline__ bool mbar_try_wait_hint_a(uint32_t bar, uint32_t parity, uint32_t hint_ns) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity), "r"(hint_ns) : "memory");
    return ok != 0;
}
// What every wait of the grouped kernels uses.  On short units two fifths of all instructions the
// first version issued were polls of warps with nothing to do (producers ahead of their ring,
// consumers ahead of their producer).  Measured alternatives, all worse for at least one kernel:
// nanosleep between polls (late wake-ups cost the decoder 1-3 points), polling without the clock
// read (faster polling hurts every kernel).
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait_a(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait_hint_a(bar, parity, 20000u))
        if (clock64() - t0 > 8000000000ll) __trap();
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// Mask keeping the n lowest bytes of a word (n <= 0: none, n >= 4: all).  PTX shifts clamp their
// amount to the register width, so no comparison is needed.
__device__ __forceinline__ uint32_t keep_low_bytes(int n) {
    uint32_t m;
    const int bits = n < 0 ? 0 : 8 * n;
    asm("shl.b32 %0, 0xFFFFFFFF, %1;" : "=r"(m) : "r"(bits));
    return ~m;
}

// Iterations cw, cw+8, ... of one stage (a tile of `len` input bytes at byte offset `sh` of
// `buf`) by consumer warp cw.
template <bool VEC16, bool SHIFTED, int kConsumerWarps, int NLUT, bool SELP>
__device__ __forceinline__ void encode_stage(const uint8_t *buf, uint32_t sh, char *dst, int len,
                                             int cw, int lane, const uint16_t *lut) {
    const uint32_t sh4 = sh & 3u, s8 = sh4 * 8;
    const uint32_t *wp = reinterpret_cast<const uint32_t *>(buf + (sh & ~3u)) + cw * 96 + lane * 3;
    char *q = dst + cw * 512 + lane * 16;
    // A stage that does not start on a word boundary needs no shifting at all: the byte
    // permutes that build the four 3-byte groups simply pick from one byte further along.
    // Group g starts at byte 3g + sh4 of the lane's four words; the selectors are uniform.
    auto sel_of = [](uint32_t o) { return (o << 12) | ((o + 1) << 8) | ((o + 2) << 4) | (o + 2); };
    const uint32_t sel0 = sel_of(sh4), sel1 = sel_of(sh4 - 1u), sel2 = sel_of(sh4 + 2u), sel3 = sel_of(sh4 + 1u);
    // Warp-iterations whose 32 lanes all hold 12 whole bytes run the plain body; the one iteration
    // that meets the end of the stage (if this warp owns it) is peeled off behind the loop, so the
    // plain body carries no end-of-unit test, and the lanes of that last iteration share ONE masked
    // encode instead of diverging into a second copy of the whole mapping.
    int ob = cw * 384;   // first byte of this warp's current iteration
    for (; ob + 384 <= len; ob += 384 * kConsumerWarps, wp += 96 * kConsumerWarps, q += 512 * kConsumerWarps) {
        uint32_t a0 = wp[0], a1 = wp[1], a2 = wp[2], a3 = 0;
        if (SHIFTED) a3 = wp[3];
        if (SHIFTED && !SELP) {
            a0 = __funnelshift_r(a0, a1, s8);
            a1 = __funnelshift_r(a1, a2, s8);
            a2 = __funnelshift_r(a2, a3, s8);
        }
        uint4 v;
        if (SHIFTED && SELP)
            v = encode_groups<NLUT>(__byte_perm(a0, a1, sel0), __byte_perm(a1, a2, sel1),
                                    __byte_perm(a1, a2, sel2), __byte_perm(a2, a3, sel3), lut);
        else
            v = encode12_mixed<NLUT>(a0, a1, a2, lut);
        if (VEC16) {
            *reinterpret_cast<uint4 *>(q) = v;
        } else {
            uint32_t *q4 = reinterpret_cast<uint32_t *>(q);
            q4[0] = v.x; q4[1] = v.y; q4[2] = v.z; q4[3] = v.w;
        }
    }
    // The iteration holding the end of the stage: every lane runs the SAME code (no divergent copy
    // of the mapping for the one lane with the end in it) -- bytes behind the end are masked off
    // with shifts that clamp, the '=' padding is patched in with selects, and only the stores
    // differ: a whole item leaves as one 16-byte store, the last item word by word.
    if (ob < len) {
        const int left = len - ob - lane * 12;   // bytes from this lane's item to the end of the stage
        const int nb = left < 0 ? 0 : (left > 12 ? 12 : left);
        uint32_t a0 = wp[0], a1 = wp[1], a2 = wp[2];
        if (SHIFTED) {
            const uint32_t a3 = wp[3];
            a0 = __funnelshift_r(a0, a1, s8);
            a1 = __funnelshift_r(a1, a2, s8);
            a2 = __funnelshift_r(a2, a3, s8);
        }
        a0 &= keep_low_bytes(nb);
        a1 &= keep_low_bytes(nb - 4);
        a2 &= keep_low_bytes(nb - 8);
        const uint4 v = encode12_mixed<(NLUT < 0 ? 0 : NLUT)>(a0, a1, a2, lut);
        // nb = 3 gi + rem: group gi is the partial one; its last one or two characters become '='
        const int gi = (nb * 11) >> 5, rem = nb - 3 * gi;
        const uint32_t pm = rem == 1 ? 0xFFFF0000u : (rem == 2 ? 0xFF000000u : 0u);
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint32_t m = g == gi ? pm : 0u;
            w[g] = (w[g] & ~m) | (0x3D3D3D3Du & m);
        }
        const int nwords = gi + (rem != 0);      // characters / 4
        if (VEC16 && nwords == 4) {
            *reinterpret_cast<uint4 *>(q) = make_uint4(w[0], w[1], w[2], w[3]);
        } else {
            uint32_t *q4 = reinterpret_cast<uint32_t *>(q);
#pragma unroll
            for (int g = 0; g < 4; ++g)
                if (g < nwords) q4[g] = w[g];
        }
    }
}

// header: [0] n_tiles, [1] next tile, [2] blocks finished (the last one re-arms [1], [2]).
// Dynamic shared memory: the ring.  RING == 0: kStages fixed slots of kStageStride bytes.
// RING > 0: a byte ring of RING bytes shared by kStages slots; every tile takes only the space
// it needs (rounded to 128 bytes), so short tiles leave room for more tiles in flight.
template <int kStages, int kConsumerWarps, int NLUT, int MIN_BLOCKS, int RING = 0, bool SELP = false>
__global__ void __launch_bounds__(32 * (1 + kConsumerWarps), MIN_BLOCKS)
b64enc_stream(const uint8_t *__restrict__ in, char *__restrict__ out,
              const TileDesc *__restrict__ desc, int *__restrict__ header,
              const int64_t *__restrict__ in_off, int64_t n_units) {
    extern __shared__ __align__(128) uint8_t ring[];
    __shared__ __align__(8) uint64_t full_bar[kStages];
    __shared__ __align__(8) uint64_t empty_bar[kStages];
    __shared__ StageMeta meta[kStages];
    __shared__ uint32_t ring_start[kStages];  // producer only: where each in-flight tile starts
    __shared__ uint16_t lut[NLUT > 0 ? 4096 : 1];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (NLUT > 0) fill_pair_lut(lut);
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kConsumerWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == 0) {
        // ------------------------------------------------------------------ producer
        const int n_tiles = header[0];
        const int64_t in_end = in_off[n_units];
        const int n_blocks = (int)gridDim.x;
        uint32_t c = 0;  // stages issued so far
        // byte-ring bookkeeping (RING > 0, lane 0 only): uses [waited, c) are in flight
        uint32_t head = 0, waited = 0;
        // first batch: static (block b takes tile b), later batches from the counter
        int base = (int)blockIdx.x, count = base < n_tiles ? 1 : 0;
        while (count > 0) {
            // this batch's descriptors, one per lane, in flight together
            int4 d0 = make_int4(0, 0, 0, 0), d1 = make_int4(0, 0, 0, 0);
            if (lane < count) {
                const int4 *dp = reinterpret_cast<const int4 *>(desc + base + lane);
                d0 = __ldg(dp);
                d1 = __ldg(dp + 1);
            }
            // the next batch (guided size), requested now and used after this one
            int nbase = 0, ncount = 0;
            if (lane == 0) {
                int rem = n_tiles - (base + count);
                int g = rem / (2 * n_blocks);
                g = g < 1 ? 1 : (g > kGrabMax ? kGrabMax : g);
                nbase = atomicAdd(&header[1], g) + n_blocks;
                ncount = nbase < n_tiles ? (n_tiles - nbase < g ? n_tiles - nbase : g) : 0;
            }
            for (int t = 0; t < count; ++t, ++c) {
                const int64_t in_pos = ((int64_t)(uint32_t)__shfl_sync(0xffffffffu, d0.y, t) << 32) |
                                       (uint32_t)__shfl_sync(0xffffffffu, d0.x, t);
                const int64_t out_pos = ((int64_t)(uint32_t)__shfl_sync(0xffffffffu, d0.w, t) << 32) |
                                        (uint32_t)__shfl_sync(0xffffffffu, d0.z, t);
                const int len = __shfl_sync(0xffffffffu, d1.x, t);
                if (lane != 0) continue;
                const uint8_t *src = in + in_pos;
                char *dst = out + out_pos;
                const uint32_t dalign = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
                // the bulk copy fetches whole 16-byte lines around the tile: keep it inside the blob
                const bool bulk_ok = in_pos >= 16 && in_pos + len + 16 <= in_end && (dalign & 3u) == 0;
                const int s = (int)(c % kStages);
                const uint32_t sh = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 15u);
                uint32_t boff = (uint32_t)s * kStageStride;
                if (RING == 0) {
                    mbar_wait(&empty_bar[s], ((c / kStages) & 1u) ^ 1u);
                } else {
                    // the slot itself: use c - kStages must be finished
                    while (waited + kStages <= c) {
                        mbar_wait(&empty_bar[waited % kStages], (waited / kStages) & 1u);
                        ++waited;
                    }
                    // space for this tile (+16 bytes the consumers may read past the end)
                    const uint32_t need = (sh + (uint32_t)len + 16u + 127u) & ~127u;
                    for (;;) {
                        if (waited == c) { head = 0; break; }              // nothing in flight
                        const uint32_t tail = ring_start[waited % kStages];  // oldest in-flight region
                        if (head > tail) {                                    // in flight: [tail, head)
                            if (head + need <= (uint32_t)RING) break;
                            if (need <= tail) { head = 0; break; }            // wrap to the front
                        } else if (head + need <= tail) {                     // in flight: [tail, ..) + [0, head)
                            break;
                        }
                        mbar_wait(&empty_bar[waited % kStages], (waited / kStages) & 1u);
                        ++waited;
                    }
                    boff = head;
                    ring_start[s] = head;
                    head += need;
                }
                StageMeta m;
                m.dst = dst;
                m.src = src;
                m.len = len;
                m.sh = sh;
                m.flags = (dalign == 0 ? 1 : 0) | (bulk_ok ? 0 : 2);
                m.pad = (int)boff;
                meta[s] = m;
                if (bulk_ok) {
                    const uint32_t bytes = (sh + (uint32_t)len + 15u) & ~15u;
                    mbar_arrive_expect_tx(&full_bar[s], bytes);
                    bulk_g2s(ring + boff, src - sh, bytes, &full_bar[s]);
                } else {
                    mbar_arrive(&full_bar[s]);
                }
            }
            base = __shfl_sync(0xffffffffu, nbase, 0);
            count = __shfl_sync(0xffffffffu, ncount, 0);
        }
        if (lane == 0) {
            // end-of-work stage, seen by every consumer warp
            const int s = (int)(c % kStages);
            if (RING == 0) {
                mbar_wait(&empty_bar[s], ((c / kStages) & 1u) ^ 1u);
            } else {
                while (waited + kStages <= c) {
                    mbar_wait(&empty_bar[waited % kStages], (waited / kStages) & 1u);
                    ++waited;
                }
            }
            meta[s].len = -1;
            mbar_arrive(&full_bar[s]);
            __threadfence();
            if (atomicAdd(&header[2], 1) == n_blocks - 1) {
                header[1] = 0;
                header[2] = 0;
                __threadfence();
            }
        }
    } else {
        // ------------------------------------------------------------------ consumers
        const int cw = warp - 1;
        for (uint32_t c = 0;; ++c) {
            const int s = (int)(c % kStages);
            mbar_wait(&full_bar[s], (c / kStages) & 1u);
            const StageMeta m = meta[s];
            if (m.len < 0) break;
            const uint8_t *buf = ring + m.pad;
            if (m.flags & 2) {
                for (int off = cw * 384 + lane * 12; off < m.len; off += 384 * kConsumerWarps)
                    encode_bytes(m.src + off, m.dst + off / 3 * 4, (m.len - off) < 12 ? (m.len - off) : 12);
            } else if (m.flags & 1) {
                if (m.sh & 3u)
                    encode_stage<true, true, kConsumerWarps, NLUT, SELP>(buf, m.sh, m.dst, m.len, cw, lane, lut);
                else
                    encode_stage<true, false, kConsumerWarps, NLUT, SELP>(buf, m.sh, m.dst, m.len, cw, lane, lut);
            } else {
                if (m.sh & 3u)
                    encode_stage<false, true, kConsumerWarps, NLUT, SELP>(buf, m.sh, m.dst, m.len, cw, lane, lut);
                else
                    encode_stage<false, false, kConsumerWarps, NLUT, SELP>(buf, m.sh, m.dst, m.len, cw, lane, lut);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);
        }
    }
}

// What the encode producer tells its consumers about a stage; the first 16 bytes are all the
// hot path reads.
struct __align__(16) EncMeta {
    char *dst;           // where this stage's characters go (solo tiles)
    int len;             // input bytes in this stage; < 0: no more work
    uint32_t info;       // bits 0-3: offset of the first input byte in its 16-byte chunk; bit 8:
                         // destination 16-byte aligned; bit 9: byte-wise stage (no bulk copy);
                         // bit 10: group tile; bits 16..: byte offset of the stage inside the ring
    int64_t in_pos;      // where the stage's first byte sits in the input blob
    int unit;            // group tile: first unit and how many
    int count;
    uint32_t offs;       // group tile: bit 2 = the units' offsets are staged behind the bytes;
                         // bit 0 / bit 1 = the in_off / out_off copy starts one entry early
    uint32_t pad_;
};

__device__ __forceinline__ uint32_t pinned_u32(uint32_t a) {
    uint32_t r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(a));   // not rematerialisable
    return r;
}

// A group tile: whole short units, one consumer warp each in turn.  Out of line: the long-unit
// path of the kernel keeps the registers and the code layout it had before group tiles existed.
template <int kConsumerWarps, int NLUT>
__device__ __noinline__ void encode_group_tile(const uint8_t *__restrict__ in, const int64_t *__restrict__ in_off,
                                               char *__restrict__ out, const int64_t *__restrict__ out_off,
                                               const uint8_t *buf, uint32_t sh, bool staged, int64_t tpos,
                                               int u0, int cnt, int len, int cw, int lane, const uint16_t *lut) {
    // offsets of 32 units at a time, one per lane: one round trip to L2 per 32 units instead of
    // three dependent ones per unit
    for (int base = 0; base < cnt; base += 32) {
        int64_t la = 0, lb = 0, lo = 0;
        if (base + lane < cnt) {
            la = in_off[u0 + base + lane];
            lb = in_off[u0 + base + lane + 1];
            lo = out_off[u0 + base + lane];
        }
        const int hi = cnt - base < 32 ? cnt - base : 32;
        for (int k = (cw + kConsumerWarps - base % kConsumerWarps) % kConsumerWarps; k < hi; k += kConsumerWarps) {
            const int64_t a = __shfl_sync(0xffffffffu, la, k);
            const int ulen = (int)(__shfl_sync(0xffffffffu, lb, k) - a);
            const int64_t uo = __shfl_sync(0xffffffffu, lo, k);
            if (ulen <= 0 || a < tpos || a + ulen > tpos + len) continue;   // empty; or offsets that are not sorted
            char *udst = out + uo;
            const uint32_t ual = (uint32_t)(reinterpret_cast<uintptr_t>(udst) & 15u);
            if (!staged || (ual & 3u)) {   // not staged, or a destination words cannot reach
                for (int off = lane * 12; off < ulen; off += 384)
                    encode_bytes(in + a + off, udst + off / 3 * 4, (ulen - off) < 12 ? (ulen - off) : 12);
                continue;
            }
            const uint32_t bo = sh + (uint32_t)(a - tpos);
            const uint8_t *ubuf = buf + (bo & ~15u);
            const uint32_t ush = bo & 15u;
            if (ual == 0) {
                if (ush & 3u) encode_stage<true, true, 1, NLUT, false>(ubuf, ush, udst, ulen, 0, lane, lut);
                else encode_stage<true, false, 1, NLUT, false>(ubuf, ush, udst, ulen, 0, lane, lut);
            } else {
                if (ush & 3u) encode_stage<false, true, 1, NLUT, false>(ubuf, ush, udst, ulen, 0, lane, lut);
                else encode_stage<false, false, 1, NLUT, false>(ubuf, ush, udst, ulen, 0, lane, lut);
            }
        }
    }
}

// encode_group_tile with the units' offsets already in shared memory (staged by the producer with
// the tile): a unit's three offsets are broadcast reads, nobody waits on a global load.
//   s_in[k] = in_off[u0 + k] (k = 0..cnt), s_out[k] = out_off[u0 + k] (k = 0..cnt-1).
template <int kConsumerWarps, int NLUT_LONG>
__device__ __noinline__ void encode_group_staged(char *__restrict__ out, const uint8_t *buf, uint32_t sh, int64_t tpos,
                                                 int cnt, int len, const int64_t *s_in, const int64_t *s_out,
                                                 int cw, int lane, const uint16_t *lut) {
    // Short units leave the load/store pipe half idle and saturate the ALU pipe (per-unit set-up is
    // all integer work), so here three of the four groups go through the pair table, not two.
    constexpr int NLUT = NLUT_LONG > 0 ? KVEDGE_GROUP_NLUT : NLUT_LONG;
    // cw: the caller rotates it from tile to tile, so that a tile of six units does not give the
    // same two warps two units every time
    for (int k = cw; k < cnt; k += kConsumerWarps) {
        const int64_t a = s_in[k];
        const int ulen = (int)(s_in[k + 1] - a);
        if (ulen <= 0 || a < tpos || a + ulen > tpos + len) continue;   // empty; or offsets that are not sorted
        char *udst = out + s_out[k];
        const uint32_t ual = (uint32_t)(reinterpret_cast<uintptr_t>(udst) & 15u);
        const uint32_t bo = sh + (uint32_t)(a - tpos);
        const uint8_t *ubuf = buf + (bo & ~15u);
        const uint32_t ush = bo & 15u;
        if (ual & 3u) {   // a destination words cannot reach: bytes, from the staged copy
            for (int off = lane * 12; off < ulen; off += 384)
                encode_bytes(buf + bo + off, udst + off / 3 * 4, (ulen - off) < 12 ? (ulen - off) : 12);
        } else if (ual == 0) {
            if (ush & 3u) encode_stage<true, true, 1, NLUT, false>(ubuf, ush, udst, ulen, 0, lane, lut);
            else encode_stage<true, false, 1, NLUT, false>(ubuf, ush, udst, ulen, 0, lane, lut);
        } else {
            if (ush & 3u) encode_stage<false, true, 1, NLUT, false>(ubuf, ush, udst, ulen, 0, lane, lut);
            else encode_stage<false, false, 1, NLUT, false>(ubuf, ush, udst, ulen, 0, lane, lut);
        }
    }
}

// One block per SM (or two) holding kGroups independent pipelines -- a producer warp, a ring
// and kConsumerWarps consumer warps each -- around ONE pair table.  Compared with
// b64enc_stream (one pipeline per block) a 2.4 KB tile occupies four consumer warps instead
// of eight, which halves the per-stage bookkeeping, and with kRing > 0 it takes only its own
// size out of the ring, so the next 11 KB tile loads while the current one is being encoded.
template <int kGroups, int kStages, int kConsumerWarps, int NLUT, int MIN_BLOCKS, int kRing>
__global__ void __launch_bounds__(32 * kGroups * (1 + kConsumerWarps), MIN_BLOCKS)
b64enc_groups(const uint8_t *__restrict__ in, char *__restrict__ out,
              const TileDesc *__restrict__ desc, int *__restrict__ header,
              const int64_t *__restrict__ in_off, int64_t n_units) {
    constexpr int kGroupWarps = 1 + kConsumerWarps;
    constexpr int kGroupBytes = kRing > 0 ? kRing : kStages * kGroupStageStride;
    static_assert(kRing == 0 || (kRing % 16 == 0 && kRing >= kStageStride), "ring must hold one full tile");
    extern __shared__ __align__(128) uint8_t dyn_smem[];
    __shared__ __align__(8) uint64_t full_bar[kGroups][kStages];
    __shared__ __align__(8) uint64_t empty_bar[kGroups][kStages];
    __shared__ __align__(16) EncMeta meta[kGroups][kStages];
    __shared__ uint32_t ring_start[kGroups][kStages];  // producer only: where each in-flight tile starts
    __shared__ uint16_t lut[NLUT > 0 ? 4096 : 1];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = warp / kGroupWarps, role = warp % kGroupWarps;
    uint8_t *ring = dyn_smem + grp * kGroupBytes;
    if (NLUT > 0) fill_pair_lut(lut);
    if (threadIdx.x == 0) {
        for (int g = 0; g < kGroups; ++g)
            for (int s = 0; s < kStages; ++s) {
                mbar_init(&full_bar[g][s], 1);
                mbar_init(&empty_bar[g][s], kConsumerWarps);
            }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    pdl_wait();   // everything above overlapped the plan; the tile table is read from here on
    const uint32_t full_a = pinned_u32(smem_u32(&full_bar[grp][0])), empty_a = pinned_u32(smem_u32(&empty_bar[grp][0]));

    const int64_t *out_off = reinterpret_cast<const int64_t *>(
        ((unsigned long long)(unsigned)header[5] << 32) | (unsigned long long)(unsigned)header[4]);
    if (role == 0) {
        // ------------------------------------------------------------------ producer
        const int n_tiles = header[0];
        const int64_t in_end = in_off[n_units];
        const int n_groups = (int)gridDim.x * kGroups;
        uint32_t c = 0;                  // tiles issued so far; tile c uses barrier pair c % kStages
        uint32_t head = 0, waited = 0;   // kRing: allocation point; tiles [waited, c) are in flight
        int base = (int)blockIdx.x * kGroups + grp, count = base < n_tiles ? 1 : 0;
        while (count > 0) {
            int4 d0 = make_int4(0, 0, 0, 0), d1 = make_int4(0, 0, 0, 0);
            if (lane < count) {
                const int4 *dp = reinterpret_cast<const int4 *>(desc + base + lane);
                d0 = __ldg(dp);
                d1 = __ldg(dp + 1);
            }
            int nbase = 0, ncount = 0;
            if (lane == 0) {
                int rem = n_tiles - (base + count);
                int g = rem / (2 * n_groups);
                g = g < 1 ? 1 : (g > kGrabMax ? kGrabMax : g);
                nbase = atomicAdd(&header[1], g) + n_groups;
                ncount = nbase < n_tiles ? (n_tiles - nbase < g ? n_tiles - nbase : g) : 0;
            }
            for (int t = 0; t < count; ++t) {
                const int64_t in_pos = ((int64_t)(uint32_t)__shfl_sync(0xffffffffu, d0.y, t) << 32) |
                                       (uint32_t)__shfl_sync(0xffffffffu, d0.x, t);
                const int64_t out_pos = ((int64_t)(uint32_t)__shfl_sync(0xffffffffu, d0.w, t) << 32) |
                                        (uint32_t)__shfl_sync(0xffffffffu, d0.z, t);
                const int len = __shfl_sync(0xffffffffu, d1.x, t);
                const int unit = __shfl_sync(0xffffffffu, d1.y, t);
                const int tflags = __shfl_sync(0xffffffffu, d1.z, t);
                const int ucount = __shfl_sync(0xffffffffu, d1.w, t);
                if (lane != 0) continue;
                if (len <= 0 || len > kStageIn) continue;   // a group of empty units; or offsets that are not sorted
                const bool is_group = (tflags & 2) != 0;
                const uint8_t *src = in + in_pos;
                char *dst = out + out_pos;
                const uint32_t dalign = is_group ? 0u : (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
                // the bulk copy fetches whole 16-byte lines around the tile: keep it inside the blob
                const bool bulk_ok = in_pos >= 16 && in_pos + len + 16 <= in_end && (dalign & 3u) == 0;
                const uint32_t s = c % kStages;
                const uint32_t sh = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 15u);
                uint32_t boff = s * kGroupStageStride;
                if (kRing == 0) {
                    mbar_wait_a(empty_a + s * 8u, ((c / kStages) & 1u) ^ 1u);
                } else {
                    while (waited + kStages <= c) {   // the barrier pair itself
                        mbar_wait_a(empty_a + (waited % kStages) * 8u, (waited / kStages) & 1u);
                        ++waited;
                    }
                    // room for this tile (+ the 16 bytes a lane may read past its last byte)
                    const uint32_t need = (sh + (uint32_t)len + 16u + 15u) & ~15u;
                    for (;;) {
                        if (waited == c) { head = 0; break; }                     // nothing in flight
                        const uint32_t tail = ring_start[grp][waited % kStages];  // oldest tile in flight
                        if (head > tail) {                                        // in flight: [tail, head)
                            if (head + need <= (uint32_t)kRing) break;
                            if (need <= tail) { head = 0; break; }                // wrap to the front
                        } else if (head + need <= tail) {                         // in flight: [tail, end) + [0, head)
                            break;
                        }
                        mbar_wait_a(empty_a + (waited % kStages) * 8u, (waited / kStages) & 1u);
                        ++waited;
                    }
                    boff = head;
                    ring_start[grp][s] = head;
                    head += need;
                }
                EncMeta m;
                m.dst = dst;
                m.len = len;
                m.info = sh | (dalign == 0 ? 0x100u : 0u) | (bulk_ok ? 0u : 0x200u) | (is_group ? 0x400u : 0u) |
                         (boff << 12);  // boff % 16 == 0
                m.in_pos = in_pos;
                m.unit = unit;
                m.count = ucount;
                m.offs = 0;
                m.pad_ = 0;
                // a group tile's unit offsets ride along (16-byte pieces of the two offset arrays
                // that cover the entries needed; never a byte outside either array)
                uint32_t bytes_i = 0, bytes_o = 0;
                const char *src_i = nullptr, *src_o = nullptr;
                if (kRing == 0 && is_group && bulk_ok && ucount <= kMaxGroupUnits) {
                    const char *pi = reinterpret_cast<const char *>(in_off + unit);
                    const char *po = reinterpret_cast<const char *>(out_off + unit);
                    const uint32_t ai = (uint32_t)(reinterpret_cast<uintptr_t>(pi) & 15u);
                    const uint32_t ao = (uint32_t)(reinterpret_cast<uintptr_t>(po) & 15u);
                    bytes_i = (ai + 8u * (uint32_t)(ucount + 1) + 15u) & ~15u;
                    bytes_o = (ao + 8u * (uint32_t)ucount + 15u) & ~15u;
                    src_i = pi - ai;
                    src_o = po - ao;
                    const char *end_i = reinterpret_cast<const char *>(in_off + n_units + 1);
                    const char *end_o = reinterpret_cast<const char *>(out_off + n_units + 1);
                    const bool fits = src_i >= reinterpret_cast<const char *>(in_off) && src_i + bytes_i <= end_i &&
                                      src_o >= reinterpret_cast<const char *>(out_off) && src_o + bytes_o <= end_o &&
                                      ((ai | ao) & 7u) == 0;
                    if (fits) m.offs = 4u | (ai >> 3) | ((ao >> 3) << 1);
                    else bytes_i = bytes_o = 0;
                }
                meta[grp][s] = m;
                if (bulk_ok) {
                    const uint32_t bytes = (sh + (uint32_t)len + 15u) & ~15u;
                    mbar_arrive_expect_tx(&full_bar[grp][s], bytes + bytes_i + bytes_o);
                    bulk_g2s(ring + boff, src - sh, bytes, &full_bar[grp][s]);
                    if (bytes_i) {
                        bulk_g2s(ring + boff + kStageStride, src_i, bytes_i, &full_bar[grp][s]);
                        bulk_g2s(ring + boff + kStageStride + kOffsPart, src_o, bytes_o, &full_bar[grp][s]);
                    }
                } else {
                    mbar_arrive(&full_bar[grp][s]);
                }
                ++c;
            }
            c = __shfl_sync(0xffffffffu, c, 0);
            base = __shfl_sync(0xffffffffu, nbase, 0);
            count = __shfl_sync(0xffffffffu, ncount, 0);
        }
        if (lane == 0) {
            // end-of-work stage, seen by every consumer warp of the group
            const uint32_t s = c % kStages;
            if (kRing == 0) {
                mbar_wait_a(empty_a + s * 8u, ((c / kStages) & 1u) ^ 1u);
            } else {
                while (waited + kStages <= c) {
                    mbar_wait_a(empty_a + (waited % kStages) * 8u, (waited / kStages) & 1u);
                    ++waited;
                }
            }
            meta[grp][s].len = -1;
            mbar_arrive(&full_bar[grp][s]);
            __threadfence();
            if (atomicAdd(&header[2], 1) == n_groups - 1) {
                header[1] = 0;
                header[2] = 0;
                __threadfence();
            }
        }
    } else {
        // ------------------------------------------------------------------ consumers
        const int cw = role - 1;
        const EncMeta *gmeta = meta[grp];
        uint32_t s = 0, ph = 0;
        int rr = 0;
        for (;;) {
            mbar_wait_a(full_a + s * 8u, ph);
            const int4 hot = *reinterpret_cast<const int4 *>(&gmeta[s]);
            const int len = hot.z;
            if (len < 0) break;
            const uint32_t info = (uint32_t)hot.w, sh = info & 15u;
            const uint8_t *buf = ring + (info >> 12);
            char *dst = reinterpret_cast<char *>(((uint64_t)(uint32_t)hot.y << 32) | (uint32_t)hot.x);
            if (info & 0x400u) {
                // units go round the consumer warps ACROSS tiles (rr: units seen so far, mod warps)
                const uint32_t offs = gmeta[s].offs;
                const int ucnt = gmeta[s].count;
                const int rcw = (cw + kConsumerWarps - rr) % kConsumerWarps;
                if (offs & 4u)
                    encode_group_staged<kConsumerWarps, NLUT>(out, buf, sh, gmeta[s].in_pos, ucnt, len,
                                                            reinterpret_cast<const int64_t *>(buf + kStageStride) + (offs & 1u),
                                                            reinterpret_cast<const int64_t *>(buf + kStageStride + kOffsPart) + ((offs >> 1) & 1u),
                                                            rcw, lane, lut);
                else
                    encode_group_tile<kConsumerWarps, NLUT>(in, in_off, out, out_off, buf, sh, (info & 0x200u) == 0, gmeta[s].in_pos,
                                                            gmeta[s].unit, ucnt, len, rcw, lane, lut);
                rr = (rr + ucnt) % kConsumerWarps;
            } else if (info & 0x200u) {
                const uint8_t *src = in + gmeta[s].in_pos;
                for (int off = cw * 384 + lane * 12; off < len; off += 384 * kConsumerWarps)
                    encode_bytes(src + off, dst + off / 3 * 4, (len - off) < 12 ? (len - off) : 12);
            } else if (info & 0x100u) {
                if (sh & 3u) encode_stage<true, true, kConsumerWarps, NLUT, false>(buf, sh, dst, len, cw, lane, lut);
                else encode_stage<true, false, kConsumerWarps, NLUT, false>(buf, sh, dst, len, cw, lane, lut);
            } else {
                if (sh & 3u) encode_stage<false, true, kConsumerWarps, NLUT, false>(buf, sh, dst, len, cw, lane, lut);
                else encode_stage<false, false, kConsumerWarps, NLUT, false>(buf, sh, dst, len, cw, lane, lut);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive_a(empty_a + s * 8u);
            if (++s == kStages) { s = 0; ph ^= 1u; }
        }
    }
}

// ----------------------------------------------------------------------------- decode, streamed
// Same producer/consumer ring for the inverse transform (what Kubernetes does to
// Secret.data.userdata before the guest reads it, _helper.tpl:64,73).  A lane takes 16
// characters and writes 12 bytes.  Every character goes through a shared-memory table that
// both validates and decodes it.  Units that are not canonical ('=' anywhere but the very
// end, CR/LF, foreign characters, length not a multiple of four) are only *flagged* here
// (status -2); b64dec_slow then re-decodes them with Go's exact semantics.
//
// The table has one entry per 7-bit character code, replicated across all 32 shared-memory
// banks (row c = 32 words): lane l always reads word l of a row, so a lookup is one wavefront
// whatever the text looks like.  (A 16384-entry pair table, one lookup per two characters,
// measured 4-5 wavefronts per lookup on real base64 text and was dropped.)  It sits at the
// start of dynamic shared memory with the ring behind it, so that the plain loop may index it
// with a whole byte: codes 128..255 read ring bytes, harmlessly, and are caught by bad_w.
__host__ __device__ constexpr int dec_stage_stride(int tile) { return (tile + 32 + 15) / 16 * 16; }
constexpr int kDecLutBytes = 128 * 32 * 4;
constexpr uint32_t kDecBadEntry = 0x01000000u;   // survives the 6- and 12-bit shifts of a group
constexpr uint32_t kDecBadMask = 0xFF000000u;

__device__ __forceinline__ void fill_dec_char_lut(uint32_t *lut) {
    for (int i = threadIdx.x; i < 128 * 32; i += blockDim.x) {
        const uint32_t c = (uint32_t)i >> 5;
        uint32_t v = kDecBadEntry;
        if (c - 65u < 26u) v = c - 65u;
        if (c - 97u < 26u) v = c - 71u;
        if (c - 48u < 10u) v = c + 4u;
        if (c == 43u) v = 62u;
        if (c == 47u) v = 63u;
        lut[i] = v;
    }
}

// General path: lane4 = 4 * lane.  bad_v collects the looked-up words (kDecBadMask bits = not
// in the alphabet), bad_w the raw characters (bit 7 of any byte = not ASCII).
__device__ __forceinline__ uint32_t decode_word_char(uint32_t w, const uint32_t *lut, uint32_t lane4,
                                                     uint32_t *bad_v, uint32_t *bad_w) {
    const char *base = reinterpret_cast<const char *>(lut);
    const uint32_t o0 = ((w << 7) & 0x3F80u) | lane4;
    const uint32_t o1 = ((w >> 1) & 0x3F80u) | lane4;
    const uint32_t o2 = ((w >> 9) & 0x3F80u) | lane4;
    const uint32_t o3 = ((w >> 17) & 0x3F80u) | lane4;
    const uint32_t v0 = *reinterpret_cast<const uint32_t *>(base + o0);
    const uint32_t v1 = *reinterpret_cast<const uint32_t *>(base + o1);
    const uint32_t v2 = *reinterpret_cast<const uint32_t *>(base + o2);
    const uint32_t v3 = *reinterpret_cast<const uint32_t *>(base + o3);
    *bad_v |= v0 | v1 | v2 | v3;
    *bad_w |= w;
    return ((v0 * 64u + v1) * 64u + v2) * 64u + v3;
}

// Plain loop: three instructions per character -- byte K of w (PRMT), row address
// byte * 128 + (table + 4 * lane) (one multiply-add), LDS.
template <int K>
__device__ __forceinline__ uint32_t dec_lookup(uint32_t w, uint32_t lut_lane) {
    uint32_t x, a, v;
    asm("prmt.b32 %0, %1, 0, %2;" : "=r"(x) : "r"(w), "n"(0x4440 + K));
    asm("mad.lo.u32 %0, %1, 128, %2;" : "=r"(a) : "r"(x), "r"(lut_lane));
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}

// Four characters -> 24 bits.  The two 12-bit halves carry a bad entry's flag up to bit 24
// or bit 30, so one OR of both halves per group collects the validity of all four.
__device__ __forceinline__ uint32_t decode_word_fast(uint32_t w, uint32_t lut_lane, uint32_t *bad_v) {
    const uint32_t v0 = dec_lookup<0>(w, lut_lane), v1 = dec_lookup<1>(w, lut_lane);
    const uint32_t v2 = dec_lookup<2>(w, lut_lane), v3 = dec_lookup<3>(w, lut_lane);
    const uint32_t p0 = v0 * 64u + v1, p1 = v2 * 64u + v3;
    *bad_v |= p0 | p1;
    return p0 * 4096u + p1;
}

// The plain iterations of one stage for one consumer warp: warp-iterations [ow, plain_end) in
// steps of kConsumerWarps, all 32 lanes full, no padding possible.  WS (0..3) is the word
// offset and BSH the byte offset of the stage's first character inside its 16-byte chunk;
// both are compile-time here so that the loop body has no selects or branches.
template <int WS, bool BSH, int kConsumerWarps>
__device__ __forceinline__ void decode_plain(const uint8_t *buf, uint32_t lut_lane, uint32_t *ow_buf,
                                             uint8_t *dst, int ow, int plain_end, uint32_t s8, int lane,
                                             uint32_t *bad_v, uint32_t *bad_w) {
    const uint4 *vp = reinterpret_cast<const uint4 *>(buf + ow + lane * 16);
    uint4 *qp = reinterpret_cast<uint4 *>(dst + ow / 4 * 3) + lane;
    const uint4 *tp = reinterpret_cast<const uint4 *>(ow_buf) + lane;
    uint32_t *sp = ow_buf + lane * 3;
    uint32_t bv = *bad_v, bw = *bad_w;
    for (; ow < plain_end; ow += 512 * kConsumerWarps, vp += 32 * kConsumerWarps, qp += 24 * kConsumerWarps) {
        const uint4 A = vp[0];
        uint4 B = make_uint4(0, 0, 0, 0);
        if (WS != 0 || BSH) B = vp[1];
        uint32_t w0, w1, w2, w3, w4;
        if (WS == 0) { w0 = A.x; w1 = A.y; w2 = A.z; w3 = A.w; w4 = B.x; }
        else if (WS == 1) { w0 = A.y; w1 = A.z; w2 = A.w; w3 = B.x; w4 = B.y; }
        else if (WS == 2) { w0 = A.z; w1 = A.w; w2 = B.x; w3 = B.y; w4 = B.z; }
        else { w0 = A.w; w1 = B.x; w2 = B.y; w3 = B.z; w4 = B.w; }
        if (BSH) {
            w0 = __funnelshift_r(w0, w1, s8);
            w1 = __funnelshift_r(w1, w2, s8);
            w2 = __funnelshift_r(w2, w3, s8);
            w3 = __funnelshift_r(w3, w4, s8);
        }
        bw |= w0 | w1 | w2 | w3;
        const uint32_t n0 = decode_word_fast(w0, lut_lane, &bv);
        const uint32_t n1 = decode_word_fast(w1, lut_lane, &bv);
        const uint32_t n2 = decode_word_fast(w2, lut_lane, &bv);
        const uint32_t n3 = decode_word_fast(w3, lut_lane, &bv);
        sp[0] = __byte_perm(n0, n1, 0x6012);
        sp[1] = __byte_perm(n1, n2, 0x5601);
        sp[2] = __byte_perm(n2, n3, 0x4560);
        __syncwarp();
        if (lane < 24) *qp = *tp;
        __syncwarp();
    }
    *bad_v = bv;
    *bad_w = bw;
}

// What the producer tells the consumers about a stage.  The first 16 bytes are all the hot
// path reads.
struct __align__(16) DecMeta {
    uint8_t *dst;        // where this stage's bytes go
    int len;             // characters in this stage; < 0: no more work
    uint32_t info;       // bits 0-3: offset of the first character in the stage buffer;
                         // bit 8: last tile of its unit; bit 9: byte-wise stage (no bulk copy);
                         // bit 10: group tile; bits 16..: byte offset of the stage inside the ring
    const uint8_t *src;  // global address of the first character (byte-wise stages, out_len)
    int unit;            // the unit (solo tile) or the first unit (group tile) ...
    int count;           // ... and how many units a group tile holds
};

// A shared address the compiler cannot rematerialise (it would otherwise rebuild the
// shared-window base around every barrier operation).
__device__ __forceinline__ uint32_t pinned_smem_u32(const void *p) {
    uint32_t a = smem_u32(p), r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(a));
    return r;
}

// The last, partial warp-iteration of a stage in the common layout (bulk-copied stage,
// 16-byte aligned destination): all lanes decode -- lanes past the end decode filler -- the
// 12-byte pieces are transposed through shared memory like in the plain loop, and the
// iteration's nb_total bytes leave as 16-byte stores plus one ragged piece.
__device__ __forceinline__ void decode_tail(const uint8_t *buf, uint32_t sh, uint32_t lut_lane, uint32_t *ow_buf,
                                            uint8_t *dst, int ow, int len, bool last_tile, int lane,
                                            uint32_t *bad_v, uint32_t *bad_w, int *npad_out) {
    const int off = ow + lane * 16;
    int nvalid = len - off;
    nvalid = nvalid < 0 ? 0 : (nvalid > 16 ? 16 : nvalid);   // a multiple of 4
    const uint4 *vp = reinterpret_cast<const uint4 *>(buf + off);
    uint4 A = make_uint4(0, 0, 0, 0), B = A;
    if (nvalid > 0) { A = vp[0]; B = vp[1]; }   // lanes past the end must not read past the tile's slot
    uint32_t w[5];
    const uint32_t ws = sh >> 2, s8 = (sh & 3u) * 8;
    if (ws == 0) { w[0] = A.x; w[1] = A.y; w[2] = A.z; w[3] = A.w; w[4] = B.x; }
    else if (ws == 1) { w[0] = A.y; w[1] = A.z; w[2] = A.w; w[3] = B.x; w[4] = B.y; }
    else if (ws == 2) { w[0] = A.z; w[1] = A.w; w[2] = B.x; w[3] = B.y; w[4] = B.z; }
    else { w[0] = A.w; w[1] = B.x; w[2] = B.y; w[3] = B.z; w[4] = B.w; }
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        w[g] = __funnelshift_r(w[g], w[g + 1], s8);
        if (g * 4 >= nvalid) w[g] = 0x41414141u;
    }
    int npad = 0;
    if (last_tile && nvalid > 0 && off + nvalid == len) {
        // '=' padding is legal only in the last quantum of the unit
        const int gq = (nvalid >> 2) - 1;
#pragma unroll
        for (int g = 0; g < 4; ++g)
            if (g == gq && (w[g] >> 24) == 0x3Du) {
                npad = 1;
                w[g] = (w[g] & 0x00FFFFFFu) | 0x41000000u;
                if (((w[g] >> 16) & 0xFFu) == 0x3Du) {
                    npad = 2;
                    w[g] = (w[g] & 0xFF00FFFFu) | 0x00410000u;
                }
            }
    }
    *bad_w |= w[0] | w[1] | w[2] | w[3];
    const uint32_t n0 = decode_word_fast(w[0], lut_lane, bad_v);
    const uint32_t n1 = decode_word_fast(w[1], lut_lane, bad_v);
    const uint32_t n2 = decode_word_fast(w[2], lut_lane, bad_v);
    const uint32_t n3 = decode_word_fast(w[3], lut_lane, bad_v);
    ow_buf[lane * 3 + 0] = __byte_perm(n0, n1, 0x6012);
    ow_buf[lane * 3 + 1] = __byte_perm(n1, n2, 0x5601);
    ow_buf[lane * 3 + 2] = __byte_perm(n2, n3, 0x4560);
    __syncwarp();
    const int owner = (len - ow - 1) >> 4;                     // lane holding the last quantum
    npad = __shfl_sync(0xffffffffu, npad, owner);
    *npad_out = npad;
    const int nb_total = (len - ow) / 4 * 3 - npad;            // 1..384
    const int r = nb_total - lane * 16;                        // bytes of this lane's 16-byte piece
    uint8_t *q = dst + ow / 4 * 3 + lane * 16;
    const uint32_t *t = ow_buf + lane * 4;
    if (lane < 24 && r >= 16) {
        *reinterpret_cast<uint4 *>(q) = *reinterpret_cast<const uint4 *>(t);
    } else if (lane < 24 && r > 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k)
            if (r >= 4 * (k + 1)) reinterpret_cast<uint32_t *>(q)[k] = t[k];
        const uint32_t lastw = t[r >> 2];
#pragma unroll
        for (int k = 0; k < 3; ++k)
            if (k < (r & 3)) q[(r & ~3) + k] = (uint8_t)(lastw >> (8 * k));
    }
    __syncwarp();
}

// The general path: byte-wise stages, destinations that are not 16-byte aligned, and any
// piece whose end needs care.  Warp-iterations ow0, ow0 + kStep, ... of a piece of `len`
// characters that starts `sh` bytes into `buf` (staged pieces) or at `src` in global memory
// (byte-wise ones).  The lane that meets the end of the unit stores its decoded length.
template <int kStep>
__device__ __forceinline__ void decode_general(const uint8_t *buf, uint32_t sh, const uint8_t *src, uint8_t *dst,
                                               int ow0, int len, bool last_tile, bool bytewise,
                                               const uint32_t *lut_words, int lane, uint32_t *bad_v, uint32_t *bad_w,
                                               int64_t *out_len_slot, int64_t chars_before) {
    const uint32_t lane4 = (uint32_t)lane * 4u;
    const uint32_t sh4 = sh & 3u, s8 = sh4 * 8;
    for (int ow = ow0; ow < len; ow += kStep) {   // warp-uniform loop
        const int off = ow + lane * 16;
        const bool active = off < len;
        const int nvalid = !active ? 0 : ((len - off) < 16 ? (len - off) : 16);  // a multiple of 4
        uint32_t w[4] = {0x41414141u, 0x41414141u, 0x41414141u, 0x41414141u};
        if (!bytewise && active) {
            // 16-byte loads (a lane-per-chunk LDS.128 is bank-conflict free; four LDS.32
            // at a 16-byte lane stride would be 4-way conflicted), then a uniform word
            // select and byte funnel for pieces that do not start on a 16-byte boundary
            const uint4 *vp = reinterpret_cast<const uint4 *>(buf + off);
            const uint4 A = vp[0];
            w[0] = A.x; w[1] = A.y; w[2] = A.z; w[3] = A.w;
            if (sh) {
                const uint4 B = vp[1];
                uint32_t w4 = B.x;
                const uint32_t ws = sh >> 2;
                if (ws == 1) { w[0] = A.y; w[1] = A.z; w[2] = A.w; w[3] = B.x; w4 = B.y; }
                else if (ws == 2) { w[0] = A.z; w[1] = A.w; w[2] = B.x; w[3] = B.y; w4 = B.z; }
                else if (ws == 3) { w[0] = A.w; w[1] = B.x; w[2] = B.y; w[3] = B.z; w4 = B.w; }
                if (sh4) {
                    w[0] = __funnelshift_r(w[0], w[1], s8);
                    w[1] = __funnelshift_r(w[1], w[2], s8);
                    w[2] = __funnelshift_r(w[2], w[3], s8);
                    w[3] = __funnelshift_r(w[3], w4, s8);
                }
            }
            if (nvalid < 16) {
#pragma unroll
                for (int g = 0; g < 4; ++g)
                    if (g * 4 >= nvalid) w[g] = 0x41414141u;
            }
        } else if (active) {
            const uint8_t *p = src + off;
#pragma unroll
            for (int i = 0; i < 16; ++i)
                if (i < nvalid) {
                    const uint32_t mk = 0xFFu << (8 * (i & 3));
                    w[i >> 2] = (w[i >> 2] & ~mk) | ((uint32_t)p[i] << (8 * (i & 3)));
                }
        }
        int npad = 0;
        const int nq = nvalid >> 2;
        if (active && last_tile && off + nvalid == len) {
            // '=' padding is legal only in the last quantum of the unit
#pragma unroll
            for (int g = 0; g < 4; ++g)
                if (g == nq - 1 && (w[g] >> 24) == 0x3Du) {
                    npad = 1;
                    w[g] = (w[g] & 0x00FFFFFFu) | 0x41000000u;
                    if (((w[g] >> 16) & 0xFFu) == 0x3Du) {
                        npad = 2;
                        w[g] = (w[g] & 0xFF00FFFFu) | 0x00410000u;
                    }
                }
            *out_len_slot = (chars_before + off + nvalid) / 4 * 3 - npad;
        }
        const uint32_t n0 = decode_word_char(w[0], lut_words, lane4, bad_v, bad_w);
        const uint32_t n1 = decode_word_char(w[1], lut_words, lane4, bad_v, bad_w);
        const uint32_t n2 = decode_word_char(w[2], lut_words, lane4, bad_v, bad_w);
        const uint32_t n3 = decode_word_char(w[3], lut_words, lane4, bad_v, bad_w);
        // 4 x 24 bits (big-endian inside each group) -> 12 bytes in stream order
        const uint32_t o0 = __byte_perm(n0, n1, 0x6012);
        const uint32_t o1 = __byte_perm(n1, n2, 0x5601);
        const uint32_t o2 = __byte_perm(n2, n3, 0x4560);
        uint8_t *q = dst + off / 4 * 3;
        const int nbytes = nq * 3 - npad;
        if (!bytewise && nbytes == 12) {
            uint32_t *q4 = reinterpret_cast<uint32_t *>(q);
            q4[0] = o0; q4[1] = o1; q4[2] = o2;
        } else if (active) {
            const uint32_t o[3] = {o0, o1, o2};
#pragma unroll
            for (int i = 0; i < 12; ++i)
                if (i < nbytes) q[i] = (uint8_t)(o[i >> 2] >> (8 * (i & 3)));
        }
    }
}

// decode_plain for a stage offset known only at run time: eight instantiations, one jump.
template <int kConsumerWarps>
__device__ __forceinline__ void decode_plain_any(uint32_t sh, const uint8_t *buf, uint32_t lut_lane, uint32_t *ow_buf,
                                                 uint8_t *dst, int ow0, int plain_end, int lane,
                                                 uint32_t *bad_v, uint32_t *bad_w) {
    const uint32_t ws = sh >> 2, s8 = (sh & 3u) * 8;
    if ((sh & 3u) == 0) {
        if (ws == 0) decode_plain<0, false, kConsumerWarps>(buf, lut_lane, ow_buf, dst, ow0, plain_end, s8, lane, bad_v, bad_w);
        else if (ws == 1) decode_plain<1, false, kConsumerWarps>(buf, lut_lane, ow_buf, dst, ow0, plain_end, s8, lane, bad_v, bad_w);
        else if (ws == 2) decode_plain<2, false, kConsumerWarps>(buf, lut_lane, ow_buf, dst, ow0, plain_end, s8, lane, bad_v, bad_w);
        else decode_plain<3, false, kConsumerWarps>(buf, lut_lane, ow_buf, dst, ow0, plain_end, s8, lane, bad_v, bad_w);
    } else {
        if (ws == 0) decode_plain<0, true, kConsumerWarps>(buf, lut_lane, ow_buf, dst, ow0, plain_end, s8, lane, bad_v, bad_w);
        else if (ws == 1) decode_plain<1, true, kConsumerWarps>(buf, lut_lane, ow_buf, dst, ow0, plain_end, s8, lane, bad_v, bad_w);
        else if (ws == 2) decode_plain<2, true, kConsumerWarps>(buf, lut_lane, ow_buf, dst, ow0, plain_end, s8, lane, bad_v, bad_w);
        else decode_plain<3, true, kConsumerWarps>(buf, lut_lane, ow_buf, dst, ow0, plain_end, s8, lane, bad_v, bad_w);
    }
}

// One block per SM holds kGroups independent pipelines -- each a producer warp, its own ring
// of kStages tiles and kConsumerWarps consumer warps -- around ONE character table.
// Dynamic shared memory: [table, kDecLutBytes][kGroups rings of kStages * stride bytes].
//
// kRing == 0: every tile gets one of kStages fixed slots of kTile (+slack) bytes.  kRing > 0:
// the group's ring is kRing bytes handed out by size, so that a 3 KB cloud-init text does not
// hold a 16 KB slot and the next big tile can load while the current one is still being
// decoded; kStages is then only the number of tiles that may be in flight.
template <int kGroups, int kStages, int kConsumerWarps, int MIN_BLOCKS, int kTile, int kRing = 0>
__global__ void __launch_bounds__(32 * kGroups * (1 + kConsumerWarps), MIN_BLOCKS)
b64dec_stream(const char *__restrict__ in, uint8_t *__restrict__ out,
              const TileDesc *__restrict__ desc, int *__restrict__ header,
              const int64_t *__restrict__ in_off, int64_t n_units,
              int64_t *__restrict__ out_len, int64_t *__restrict__ status) {
    constexpr int kGroupWarps = 1 + kConsumerWarps;
    constexpr int kDecStageStride = dec_stage_stride(kTile);
    constexpr int kGroupBytes = kRing > 0 ? kRing : kStages * kDecStageStride;
    static_assert(kGroupBytes >= kDecLutBytes, "byte-indexed lookups must stay inside the block's shared memory");
    static_assert(kTile % 512 == 0, "a stage is a whole number of warp-iterations");
    static_assert(kRing == 0 || (kRing % 16 == 0 && kRing >= kTile + 64), "ring must hold one full tile");
    extern __shared__ __align__(128) uint8_t dyn_smem[];
    __shared__ __align__(8) uint64_t full_bar[kGroups][kStages];
    __shared__ __align__(8) uint64_t empty_bar[kGroups][kStages];
    __shared__ DecMeta meta[kGroups][kStages];
    __shared__ uint32_t ring_start[kGroups][kStages];  // producer only: where each in-flight tile starts
    __shared__ __align__(16) uint32_t obuf[kGroups * kConsumerWarps][96];  // per-warp store transpose
    uint32_t *lut_words = reinterpret_cast<uint32_t *>(dyn_smem);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = warp / kGroupWarps, role = warp % kGroupWarps;
    uint8_t *ring = dyn_smem + kDecLutBytes + grp * kGroupBytes;
    fill_dec_char_lut(lut_words);
    if (threadIdx.x == 0) {
        for (int g = 0; g < kGroups; ++g)
            for (int s = 0; s < kStages; ++s) {
                mbar_init(&full_bar[g][s], 1);
                mbar_init(&empty_bar[g][s], kConsumerWarps);
            }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    pdl_wait();   // everything above overlapped the plan; the tile table is read from here on
    const uint32_t full_a = pinned_smem_u32(&full_bar[grp][0]), empty_a = pinned_smem_u32(&empty_bar[grp][0]);

    if (role == 0) {
        // ------------------------------------------------------------------ producer
        const int n_tiles = header[0];
        const int64_t in_end = in_off[n_units];
        const int n_groups = (int)gridDim.x * kGroups;
        uint32_t c = 0;                  // tiles issued so far; tile c uses barrier pair c % kStages
        uint32_t head = 0, waited = 0;   // kRing: allocation point; tiles [waited, c) are in flight
        int base = (int)blockIdx.x * kGroups + grp, count = base < n_tiles ? 1 : 0;
        while (count > 0) {
            int4 d0 = make_int4(0, 0, 0, 0), d1 = make_int4(0, 0, 0, 0);
            if (lane < count) {
                const int4 *dp = reinterpret_cast<const int4 *>(desc + base + lane);
                d0 = __ldg(dp);
                d1 = __ldg(dp + 1);
            }
            int nbase = 0, ncount = 0;
            if (lane == 0) {
                int rem = n_tiles - (base + count);
                int g = rem / (2 * n_groups);
                g = g < 1 ? 1 : (g > kGrabMax ? kGrabMax : g);
                nbase = atomicAdd(&header[1], g) + n_groups;
                ncount = nbase < n_tiles ? (n_tiles - nbase < g ? n_tiles - nbase : g) : 0;
            }
            for (int t = 0; t < count; ++t) {
                const int64_t in_pos = ((int64_t)(uint32_t)__shfl_sync(0xffffffffu, d0.y, t) << 32) |
                                       (uint32_t)__shfl_sync(0xffffffffu, d0.x, t);
                const int64_t out_pos = ((int64_t)(uint32_t)__shfl_sync(0xffffffffu, d0.w, t) << 32) |
                                        (uint32_t)__shfl_sync(0xffffffffu, d0.z, t);
                const int len = __shfl_sync(0xffffffffu, d1.x, t);
                const int unit = __shfl_sync(0xffffffffu, d1.y, t);
                const int tflags = __shfl_sync(0xffffffffu, d1.z, t);
                const int ucount = __shfl_sync(0xffffffffu, d1.w, t);
                if (lane != 0) continue;
                const bool is_group = (tflags & 2) != 0;
                if (!is_group && (len & 3)) continue;  // unit length not a multiple of 4: flagged by the plan (-2)
                if (len <= 0 || len > kTile) continue;  // a group of empty units; or offsets that are not sorted
                const uint8_t *src = reinterpret_cast<const uint8_t *>(in) + in_pos;
                uint8_t *dst = out + out_pos;
                const bool dst4 = is_group || (reinterpret_cast<uintptr_t>(dst) & 3u) == 0;
                const bool bulk_ok = in_pos >= 16 && in_pos + len + 16 <= in_end && dst4;
                const uint32_t s = c % kStages;
                const uint32_t sh = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 15u);
                uint32_t boff = s * kDecStageStride;
                if (kRing == 0) {
                    mbar_wait_a(empty_a + s * 8u, ((c / kStages) & 1u) ^ 1u);
                } else {
                    // the barrier pair itself: tile c - kStages must be finished
                    while (waited + kStages <= c) {
                        mbar_wait_a(empty_a + (waited % kStages) * 8u, (waited / kStages) & 1u);
                        ++waited;
                    }
                    // room for this tile (+ the 16-byte chunk a lane may read past its last character)
                    const uint32_t need = (sh + (uint32_t)len + 32u + 15u) & ~15u;
                    for (;;) {
                        if (waited == c) { head = 0; break; }                     // nothing in flight
                        const uint32_t tail = ring_start[grp][waited % kStages];  // oldest tile in flight
                        if (head > tail) {                                        // in flight: [tail, head)
                            if (head + need <= (uint32_t)kRing) break;
                            if (need <= tail) { head = 0; break; }                // wrap to the front
                        } else if (head + need <= tail) {                         // in flight: [tail, end) + [0, head)
                            break;
                        }
                        mbar_wait_a(empty_a + (waited % kStages) * 8u, (waited / kStages) & 1u);
                        ++waited;
                    }
                    boff = head;
                    ring_start[grp][s] = head;
                    head += need;
                }
                DecMeta m;
                m.dst = dst;
                m.len = len;
                m.info = sh | ((uint32_t)(tflags & 1) << 8) | (bulk_ok ? 0u : 0x200u) | (is_group ? 0x400u : 0u) |
                         (boff << 12);  // boff is a multiple of 16
                m.src = src;
                m.unit = unit;
                m.count = ucount;
                meta[grp][s] = m;
                if (bulk_ok) {
                    const uint32_t bytes = (sh + (uint32_t)len + 15u) & ~15u;
                    mbar_arrive_expect_tx(&full_bar[grp][s], bytes);
                    bulk_g2s(ring + boff, src - sh, bytes, &full_bar[grp][s]);
                } else {
                    mbar_arrive(&full_bar[grp][s]);
                }
                ++c;
            }
            c = __shfl_sync(0xffffffffu, c, 0);
            base = __shfl_sync(0xffffffffu, nbase, 0);
            count = __shfl_sync(0xffffffffu, ncount, 0);
        }
        if (lane == 0) {
            const uint32_t s = c % kStages;
            if (kRing == 0) {
                mbar_wait_a(empty_a + s * 8u, ((c / kStages) & 1u) ^ 1u);
            } else {
                while (waited + kStages <= c) {
                    mbar_wait_a(empty_a + (waited % kStages) * 8u, (waited / kStages) & 1u);
                    ++waited;
                }
            }
            meta[grp][s].len = -1;
            mbar_arrive(&full_bar[grp][s]);
            __threadfence();
            if (atomicAdd(&header[2], 1) == n_groups - 1) {
                header[1] = 0;
                header[2] = 0;
                __threadfence();
            }
        }
    } else {
        // ------------------------------------------------------------------ consumers
        const int cw = role - 1;
        const uint32_t lane4 = (uint32_t)lane * 4u;
        const uint32_t lut_lane = pinned_smem_u32(lut_words) + lane4;
        uint32_t *ow_buf = obuf[grp * kConsumerWarps + cw];
        const DecMeta *gmeta = meta[grp];
        const int64_t *out_off = reinterpret_cast<const int64_t *>(
            ((unsigned long long)(unsigned)header[5] << 32) | (unsigned long long)(unsigned)header[4]);
        uint32_t s = 0, ph = 0;
        for (;;) {
            mbar_wait_a(full_a + s * 8u, ph);
            const int4 hot = *reinterpret_cast<const int4 *>(&gmeta[s]);
            const int len = hot.z;
            if (len < 0) break;
            const uint32_t info = (uint32_t)hot.w, sh = info & 15u;
            const bool last_tile = (info & 0x100u) != 0, bytewise = (info & 0x200u) != 0;
            const uint8_t *buf = ring + ((info >> 12) & ~15u);
            uint8_t *dst = reinterpret_cast<uint8_t *>(((uint64_t)(uint32_t)hot.y << 32) | (uint32_t)hot.x);
            if (info & 0x400u) {
                // group tile: whole short units, one warp each in turn
                const int u0 = gmeta[s].unit, cnt = gmeta[s].count;
                const uint8_t *tsrc = gmeta[s].src;
                const int64_t tpos = tsrc - reinterpret_cast<const uint8_t *>(in);
                for (int j = cw; j < cnt; j += kConsumerWarps) {
                    const int u = u0 + j;
                    const int64_t a = in_off[u];
                    const int ulen = (int)(in_off[u + 1] - a);
                    if (ulen <= 0 || (ulen & 3)) continue;   // empty, or already flagged by the plan
                    if (a < tpos || a + ulen > tpos + len) continue;   // offsets that are not sorted
                    uint8_t *udst = out + out_off[u];
                    const uint32_t bo = sh + (uint32_t)(a - tpos);
                    const uint8_t *ubuf = buf + (bo & ~15u);
                    const uint32_t ush = bo & 15u;
                    const uint32_t ual = (uint32_t)(reinterpret_cast<uintptr_t>(udst) & 15u);
                    uint32_t bv = 0, bw = 0;
                    if (!bytewise && ual == 0) {
                        const int plain_end = (ulen - 1) / 512 * 512;   // iterations that cannot hold the end
                        if (plain_end > 0) decode_plain_any<1>(ush, ubuf, lut_lane, ow_buf, udst, 0, plain_end, lane, &bv, &bw);
                        int npad = 0;
                        decode_tail(ubuf, ush, lut_lane, ow_buf, udst, plain_end, ulen, true, lane, &bv, &bw, &npad);
                        if (lane == 0) out_len[u] = (int64_t)(ulen / 4 * 3 - npad);
                    } else {
                        decode_general<512>(ubuf, ush, tsrc + (a - tpos), udst, 0, ulen, true, bytewise || (ual & 3u) != 0,
                                            lut_words, lane, &bv, &bw, &out_len[u], 0);
                    }
                    if ((bv & kDecBadMask) || (bw & 0x80808080u)) status[u] = -2;
                }
            } else {
                uint32_t bad_v = 0, bad_w = 0;
                const bool dst16 = (reinterpret_cast<uintptr_t>(dst) & 15u) == 0;
                // Plain iterations (all 32 lanes full, the unit cannot end inside) go through the
                // branch-free templated loop; the stage's last warp-iteration takes decode_tail;
                // byte-wise stages and unaligned destinations take the general code.
                int ow0 = cw * 512;
                if (!bytewise && dst16) {
                    const int plain_len = last_tile ? len - 1 : len;          // iterations with ow + 512 <= plain_len
                    const int n_it = plain_len >= ow0 + 512 ? (int)((uint32_t)(plain_len - ow0 - 512) / (512u * kConsumerWarps)) + 1 : 0;
                    const int plain_end = ow0 + n_it * 512 * kConsumerWarps;  // first ow not handled here
                    if (n_it > 0) {
                        decode_plain_any<kConsumerWarps>(sh, buf, lut_lane, ow_buf, dst, ow0, plain_end, lane, &bad_v, &bad_w);
                        ow0 = plain_end;
                    }
                    if (ow0 < len) {   // this warp owns the stage's last warp-iteration
                        int npad = 0;
                        decode_tail(buf, sh, lut_lane, ow_buf, dst, ow0, len, last_tile, lane, &bad_v, &bad_w, &npad);
                        if (last_tile && lane == 0) {
                            const int unit = gmeta[s].unit;
                            out_len[unit] = (int64_t)(gmeta[s].src - reinterpret_cast<const uint8_t *>(in) - in_off[unit] + len) / 4 * 3 - npad;
                        }
                        ow0 = len;
                    }
                }
                if (ow0 < len) {
                    const int unit = gmeta[s].unit;
                    const uint8_t *src = gmeta[s].src;
                    decode_general<512 * kConsumerWarps>(buf, sh, src, dst, ow0, len, last_tile, bytewise, lut_words, lane,
                                                         &bad_v, &bad_w, &out_len[unit],
                                                         (int64_t)(src - reinterpret_cast<const uint8_t *>(in)) - in_off[unit]);
                }
                if ((bad_v & kDecBadMask) || (bad_w & 0x80808080u)) status[gmeta[s].unit] = -2;  // benign race: every writer stores the same value
            }
            __syncwarp();
            if (lane == 0) mbar_arrive_a(empty_a + s * 8u);
            if (++s == kStages) { s = 0; ph ^= 1u; }
        }
    }
}

}  // namespace stream

// ----------------------------------------------------------------------------- decode
// One alphabet character -> its 6-bit value, or a value >= 64 if it is not in the
// alphabet.
__device__ __forceinline__ uint32_t dec6(uint32_t c) {
    uint32_t v = 0x80u;
    if (c - 65u < 26u) v = c - 65u;
    if (c - 97u < 26u) v = c - 71u;
    if (c - 48u < 10u) v = c + 4u;
    if (c == 43u) v = 62u;
    if (c == 47u) v = 63u;
    return v;
}

// Sequential decode of the units the fast path flagged (status == -2): '\r'/'\n'
// inside the text, padding in odd places, characters outside the alphabet, lengths
// that are not a multiple of four.  Follows Go's encoding/base64 decodeQuantum
// (padded, non-strict) step by step so that both the bytes and the reported
// CorruptInputError offset are Go's.  One thread per flagged unit: this is the
// cold path (Kubernetes Secrets are canonical).
__device__ __forceinline__ bool is_nl(uint8_t c) { return c == '\n' || c == '\r'; }

__global__ void b64dec_slow(const char *__restrict__ in, const int64_t *__restrict__ in_off,
                            int64_t n_units, uint8_t *__restrict__ out,
                            const int64_t *__restrict__ out_off, int64_t *__restrict__ out_len,
                            int64_t *__restrict__ status) {
    for (int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u < n_units;
         u += (int64_t)gridDim.x * blockDim.x) {
        if (status[u] != -2) continue;
        const uint8_t *src = reinterpret_cast<const uint8_t *>(in) + in_off[u];
        const int64_t len = in_off[u + 1] - in_off[u];
        uint8_t *dst = out + out_off[u];
        int64_t si = 0, n = 0, err = -1;
        bool done = false;
        while (!done) {
            uint32_t dbuf[4] = {0, 0, 0, 0};
            int dlen = 4;
            bool end = false;
            for (int j = 0; j < 4; ++j) {
                if (si == len) {
                    if (j != 0) err = si - j;
                    dlen = 0;
                    done = true;
                    break;
                }
                uint8_t c = src[si++];
                uint32_t v = dec6(c);
                if (v < 64u) { dbuf[j] = v; continue; }
                if (is_nl(c)) { --j; continue; }
                if (c != '=') { err = si - 1; dlen = 0; done = true; break; }
                if (j < 2) { err = si - 1; dlen = 0; done = true; break; }
                if (j == 2) {
                    while (si < len && is_nl(src[si])) ++si;
                    if (si == len) { err = len; dlen = 0; done = true; break; }
                    if (src[si] != '=') { err = si - 1; dlen = 0; done = true; break; }
                    ++si;
                }
                while (si < len && is_nl(src[si])) ++si;
                if (si < len) err = si;  // trailing garbage (bytes of this quantum are kept)
                dlen = j;
                end = true;
                break;
            }
            if (dlen >= 2) {
                uint32_t val = (dbuf[0] << 18) | (dbuf[1] << 12) | (dbuf[2] << 6) | dbuf[3];
                dst[n] = (uint8_t)(val >> 16);
                if (dlen >= 3) dst[n + 1] = (uint8_t)(val >> 8);
                if (dlen >= 4) dst[n + 2] = (uint8_t)val;
                n += dlen - 1;
            }
            if (end) done = true;
        }
        out_len[u] = n;
        status[u] = err;  // -1 when the unit decoded cleanly
    }
}

// ----------------------------------------------------------------------------- host-side plan
// The same tile table as plan_count + plan_emit, built on the host.  The host-buffer entry
// points know every offset already, so they write the table into pinned memory and ship it
// with the offsets in one copy instead of launching two planning kernels per chunk.
int64_t plan_tiles_host(const int64_t *in_off, const int64_t *out_off, int64_t n_units, int tile_in,
                        int small_max, int mode, const ScratchLayout &L, void *scratch_host,
                        const int64_t *d_out_off, int64_t *out_len, int64_t *status) {
    char *base = static_cast<char *>(scratch_host);
    int *header = reinterpret_cast<int *>(base);
    TileDesc *desc = reinterpret_cast<TileDesc *>(base + L.desc_off);
    int64_t t = 0;
    for (int64_t u = 0; u < n_units; ++u) {
        const int64_t ibeg = in_off[u];
        int64_t len = in_off[u + 1] - ibeg;
        if (len < 0) len = 0;
        if (mode == 1) {
            out_len[u] = len / 4 * 3;
            status[u] = (len & 3) ? -2 : -1;
        }
        if (len <= small_max) {   // same rule as plan_count / plan_emit
            if (small_max > 0 && group_starts_at(in_off, u, small_max, tile_in - small_max)) {
                int64_t last = u + 1;
                while (last < n_units && in_off[last + 1] - in_off[last] <= small_max &&
                       !group_starts_at(in_off, last, small_max, tile_in - small_max))
                    ++last;
                TileDesc d;
                d.in_pos = ibeg;
                d.unit = (int32_t)u;
                if (last - u == 1) {
                    d.out_pos = out_off[u];
                    d.len = (int32_t)len;
                    d.flags = 1;
                    d.pad = 0;
                } else {
                    d.out_pos = 0;
                    d.len = (int32_t)(in_off[last] - ibeg);
                    d.flags = 2;
                    d.pad = (int32_t)(last - u);
                }
                desc[t++] = d;
            }
            continue;
        }
        for (int64_t s = 0; s < len; s += tile_in, ++t) {
            TileDesc d;
            d.in_pos = ibeg + s;
            d.out_pos = out_off[u] + (mode == 0 ? s / 3 * 4 : s / 4 * 3);
            const int64_t rem = len - s;
            d.len = (int32_t)(rem < tile_in ? rem : tile_in);
            d.unit = (int32_t)u;
            d.flags = (s + tile_in >= len) ? 1 : 0;
            d.pad = 0;
            desc[t] = d;
        }
    }
    header[0] = (int)t;
    header[1] = 0;
    header[2] = 0;
    header[3] = 0;
    const unsigned long long op = (unsigned long long)reinterpret_cast<uintptr_t>(d_out_off);
    header[4] = (int)(unsigned)(op & 0xffffffffull);
    header[5] = (int)(unsigned)(op >> 32);
    return t;
}

// ----------------------------------------------------------------------------- launchers
// Per (device, kernel): the dynamic-shared-memory opt-in (a per-device function attribute) done
// and the resident blocks per SM known.  The *_dev entry points may be called from any thread,
// and a process may hold unlike GPUs, hence the lock and the device in the key.
namespace {
std::mutex g_launch_mu;
std::map<std::pair<int, const void *>, int> g_kernel_occ;
}  // namespace

static cudaError_t kernel_ready(const void *fn, int threads, size_t dyn, int *occ_out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(g_launch_mu);
    auto it = g_kernel_occ.find({dev, fn});
    if (it == g_kernel_occ.end()) {
        e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
        if (e != cudaSuccess) return e;
        int occ = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, threads, dyn);
        if (e != cudaSuccess) return e;
        it = g_kernel_occ.emplace(std::make_pair(dev, fn), occ > 0 ? occ : 1).first;
    }
    *occ_out = it->second;
    return cudaSuccess;
}

void reset_launch_caches() {
    std::lock_guard<std::mutex> lk(g_launch_mu);
    g_kernel_occ.clear();
}

static cudaError_t plan(const int64_t *d_in_off, const int64_t *d_out_off, int64_t n_units,
                        int tile_in, int small_max, int mode, const ScratchLayout &L, void *d_scratch,
                        int64_t *d_out_len, int64_t *d_status, cudaStream_t stream) {
    char *base = static_cast<char *>(d_scratch);
    int *header = reinterpret_cast<int *>(base);
    int *block_sums = reinterpret_cast<int *>(base + L.block_sums_off);
    TileDesc *desc = reinterpret_cast<TileDesc *>(base + L.desc_off);
    unsigned nb = (unsigned)L.n_plan_blocks;
    plan_count<<<nb, kPlanBlock, 0, stream>>>(d_in_off, n_units, tile_in, small_max, block_sums);
    // more than ~2048 tiles per plan block on average: a few very long units; share them out
    int64_t ny = L.max_tiles / ((int64_t)nb * 2048);
    ny = ny < 1 ? 1 : (ny > 32 ? 32 : ny);
    const dim3 grid(nb, (unsigned)ny);
    if (mode == 0)
        return launch_after_kernel(plan_emit<0>, grid, dim3(kPlanBlock), 0, stream, d_in_off, d_out_off, n_units, tile_in,
                                   small_max, (const int *)block_sums, header, desc, (int)L.max_tiles,
                                   (int64_t *)nullptr, (int64_t *)nullptr);
    return launch_after_kernel(plan_emit<1>, grid, dim3(kPlanBlock), 0, stream, d_in_off, d_out_off, n_units, tile_in,
                               small_max, (const int *)block_sums, header, desc, (int)L.max_tiles, d_out_len, d_status);
}

// The encode kernels' shapes.  Entry 0 of each table is the product; the rest are sweep shapes
// of the tools build (tools/sweep_enc.py), absent from the product library.
using EncKernel = void (*)(const uint8_t *, char *, const TileDesc *, int *, const int64_t *, int64_t);
struct EncGroupVariant {
    EncKernel fn; int groups, stages, consumers, ring;
    int threads() const { return 32 * groups * (1 + consumers); }
    size_t dyn() const { return (size_t)groups * (ring ? (size_t)ring : (size_t)stages * stream::kGroupStageStride); }
};
static const EncGroupVariant kEncGroup[] = {
    {stream::b64enc_groups<5, 3, 4, 2, 1, 0>, 5, 3, 4, 0},   // product
#ifdef KVEDGE_TOOLS
    {stream::b64enc_groups<5, 3, 4, 2, 1, 0>, 5, 3, 4, 0},   // (201 was 6 x 3 x 4: no longer fits since stages carry unit offsets)
    {stream::b64enc_groups<6, 8, 4, 2, 1, 36480>, 6, 8, 4, 36480},
    {stream::b64enc_groups<5, 3, 4, 3, 1, 0>, 5, 3, 4, 0},
    {stream::b64enc_groups<4, 3, 4, 2, 1, 0>, 4, 3, 4, 0},
    {stream::b64enc_groups<4, 4, 4, 2, 1, 0>, 4, 4, 4, 0},
    {stream::b64enc_groups<4, 3, 5, 2, 1, 0>, 4, 3, 5, 0},
    {stream::b64enc_groups<5, 3, 4, 2, 1, 0>, 5, 3, 4, 0},   // (207 was 5 x 4 x 4: likewise)
    {stream::b64enc_groups<5, 3, 3, 2, 1, 0>, 5, 3, 3, 0},
#endif
};
struct EncStreamVariant {
    EncKernel fn; int stages, consumers, ring;
    int threads() const { return 32 * (1 + consumers); }
    size_t dyn() const { return ring ? (size_t)ring : (size_t)stages * stream::kStageStride; }
};
static const EncStreamVariant kEncStream[] = {
    {stream::b64enc_stream<3, 8, 0, 4>, 3, 8, 0},            // product, tiny batches: all-arithmetic, no table to fill
#ifdef KVEDGE_TOOLS
    {stream::b64enc_stream<4, 8, 2, 3>, 4, 8, 0},
    {stream::b64enc_stream<3, 8, 1, 4>, 3, 8, 0},  {stream::b64enc_stream<3, 8, 3, 4>, 3, 8, 0},
    {stream::b64enc_stream<3, 8, 2, 4>, 3, 8, 0},  {stream::b64enc_stream<3, 12, 2, 3>, 3, 12, 0},
    {stream::b64enc_stream<2, 8, 2, 5>, 2, 8, 0},  {stream::b64enc_stream<4, 12, 2, 3>, 4, 12, 0},
    {stream::b64enc_stream<3, 16, 2, 2>, 3, 16, 0}, {stream::b64enc_stream<6, 16, 2, 2>, 6, 16, 0},
    {stream::b64enc_stream<8, 8, 2, 4, 36352>, 8, 8, 36352}, {stream::b64enc_stream<6, 8, 2, 4, 36352>, 6, 8, 36352},
    {stream::b64enc_stream<12, 8, 2, 4, 36352>, 12, 8, 36352}, {stream::b64enc_stream<8, 8, 2, 3, 57344>, 8, 8, 57344},
    {stream::b64enc_stream<3, 8, -1, 4>, 3, 8, 0},           // 14: traffic-only skeleton (wrong bytes; tools build only)
    {stream::b64enc_stream<3, 8, 2, 4, 0, true>, 3, 8, 0}, {stream::b64enc_stream<3, 8, 4, 4>, 3, 8, 0},
    {stream::b64enc_stream<3, 8, 2, 4>, 3, 8, 0},            // 17: the round-1 one-pipeline kernel
#endif
};

// Which encode kernel a batch gets: the grouped kernel (200) unless the batch is tiny (0).  The
// plan and the run must agree on it, because only the grouped kernels understand group tiles.
// The product library knows these two and nothing else; the KVEDGE_ENC_VARIANT knob and the
// other shapes exist only in the tools build (-DKVEDGE_TOOLS, libkvedge_b200_tools.so), so the
// bytes a caller gets can never depend on its environment.
static int enc_variant_for(const ScratchLayout &L) {
    int latched = 0;
#ifdef KVEDGE_TOOLS
    static const int env_variant = [] {
        const char *v = getenv("KVEDGE_ENC_VARIANT");
        const int x = v ? atoi(v) : 0;
        return x < 0 ? 0 : x;
    }();
    latched = env_variant;
#endif
    return latched == 0 && L.max_tiles > 32 ? 200 : latched;
}
int enc_small_max(const ScratchLayout &L) { return enc_variant_for(L) >= 200 ? kEncTileIn / 4 : 0; }

cudaError_t launch_b64enc_plan(const int64_t *d_in_off, int64_t n_units, int64_t total_in_bytes,
                               const int64_t *d_out_off, void *d_scratch, size_t scratch_bytes,
                               cudaStream_t stream) {
    if (n_units <= 0) return cudaSuccess;
    ScratchLayout L = scratch_layout(n_units, total_in_bytes, kEncTileIn);
    if (scratch_bytes < L.total) return cudaErrorInvalidValue;
    return plan(d_in_off, d_out_off, n_units, kEncTileIn, enc_small_max(L), 0, L, d_scratch, nullptr, nullptr, stream);
}

cudaError_t launch_b64enc_run(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                              int64_t total_in_bytes, char *d_out, void *d_scratch,
                              size_t scratch_bytes, cudaStream_t stream, int sm_count) {
    if (n_units <= 0) return cudaSuccess;
    ScratchLayout L = scratch_layout(n_units, total_in_bytes, kEncTileIn);
    if (scratch_bytes < L.total) return cudaErrorInvalidValue;
    const int variant = enc_variant_for(L);
    char *base = static_cast<char *>(d_scratch);
    const TileDesc *desc = reinterpret_cast<const TileDesc *>(base + L.desc_off);
    int *header = reinterpret_cast<int *>(base);
    if (variant >= 200) {  // grouped pipelines: several producer/consumer groups per block around one table
        const int ng = (int)(sizeof kEncGroup / sizeof *kEncGroup);
        const EncGroupVariant &gv = kEncGroup[variant - 200 < ng ? variant - 200 : 0];
        int occ = 1;
        cudaError_t e = kernel_ready(reinterpret_cast<const void *>(gv.fn), gv.threads(), gv.dyn(), &occ);
        if (e != cudaSuccess) return e;
        int64_t grid = (int64_t)sm_count * occ;
        if (grid > (L.max_tiles + gv.groups - 1) / gv.groups) grid = (L.max_tiles + gv.groups - 1) / gv.groups;
        if (grid < 1) grid = 1;
        return launch_after_kernel(gv.fn, dim3((unsigned)grid), dim3(gv.threads()), gv.dyn(), stream, d_in, d_out, desc,
                                   header, d_in_off, n_units);
    }
    // one pipeline per block: tiny batches (a single render) take the all-arithmetic build, which
    // has no table to fill
    const int nv = (int)(sizeof kEncStream / sizeof *kEncStream);
    const EncStreamVariant &sv = kEncStream[variant < nv ? variant : 0];
    int occ = 1;
    cudaError_t e = kernel_ready(reinterpret_cast<const void *>(sv.fn), sv.threads(), sv.dyn(), &occ);
    if (e != cudaSuccess) return e;
    int64_t grid = (int64_t)sm_count * occ;
    if (grid > L.max_tiles) grid = L.max_tiles;
    if (grid < 1) grid = 1;
    sv.fn<<<(unsigned)grid, sv.threads(), sv.dyn(), stream>>>(d_in, d_out, desc, header, d_in_off, n_units);
    return cudaGetLastError();
}

// Opt-ins and occupancy queries of every kernel the product may launch, done now for the current
// device (kv_b64enc_batch_dev_static's set-up: nothing of the kind is left for a later call).
cudaError_t warm_b64enc_kernels() {
    int occ = 0;
    cudaError_t e = kernel_ready(reinterpret_cast<const void *>(kEncGroup[0].fn), kEncGroup[0].threads(), kEncGroup[0].dyn(), &occ);
    if (e != cudaSuccess) return e;
    return kernel_ready(reinterpret_cast<const void *>(kEncStream[0].fn), kEncStream[0].threads(), kEncStream[0].dyn(), &occ);
}

cudaError_t launch_b64enc(const uint8_t *d_in, const int64_t *d_in_off, int64_t n_units,
                          int64_t total_in_bytes, char *d_out, const int64_t *d_out_off,
                          void *d_scratch, size_t scratch_bytes, cudaStream_t stream,
                          int sm_count) {
    cudaError_t e = launch_b64enc_plan(d_in_off, n_units, total_in_bytes, d_out_off, d_scratch,
                                       scratch_bytes, stream);
    if (e != cudaSuccess) return e;
    return launch_b64enc_run(d_in, d_in_off, n_units, total_in_bytes, d_out, d_scratch,
                             scratch_bytes, stream, sm_count);
}

// The streamed decode kernel's shapes.  Entry 0 is the product; the others exist only in the
// tools build (tools/sweep_dec.py, KVEDGE_DEC_VARIANT=n).
using DecKernel = void (*)(const char *, uint8_t *, const TileDesc *, int *, const int64_t *, int64_t,
                           int64_t *, int64_t *);
struct DecVariant {
    DecKernel fn; int groups, stages, consumers, tile, ring;
    int threads() const { return 32 * groups * (1 + consumers); }
    size_t dyn() const {
        return (size_t)stream::kDecLutBytes + (size_t)groups * (ring ? (size_t)ring : (size_t)stages * stream::dec_stage_stride(tile));
    }
};
static const DecVariant kDec[] = {
    {stream::b64dec_stream<5, 8, 5, 1, 16384, 40448>, 5, 8, 5, 16384, 40448},  // product
#ifdef KVEDGE_TOOLS
    {stream::b64dec_stream<1, 3, 8, 4, 16384>, 1, 3, 8, 16384, 0},             // one pipeline per block, 3 blocks/SM
    {stream::b64dec_stream<6, 2, 4, 1, 16384>, 6, 2, 4, 16384, 0},             // grouped, fixed slots
    {stream::b64dec_stream<5, 4, 5, 1, 16384, 40960>, 5, 4, 5, 16384, 40960},
    {stream::b64dec_stream<6, 4, 4, 1, 16384, 34256>, 6, 4, 4, 16384, 34256},
    {stream::b64dec_stream<5, 8, 4, 1, 16384, 40448>, 5, 8, 4, 16384, 40448},
    {stream::b64dec_stream<4, 8, 6, 1, 16384, 50688>, 4, 8, 6, 16384, 50688},
#endif
};

static int dec_variant() {
#ifdef KVEDGE_TOOLS
    static const int v = [] {
        const char *e = getenv("KVEDGE_DEC_VARIANT");
        const int x = e ? atoi(e) : 0;
        return x < 0 || x >= (int)(sizeof kDec / sizeof *kDec) ? 0 : x;
    }();
    return v;
#else
    return 0;
#endif
}

int dec_tile_in() { return kDec[dec_variant()].tile; }
int dec_small_max() { return dec_tile_in() / 4; }

cudaError_t launch_b64dec(const char *d_in, const int64_t *d_in_off, int64_t n_units,
                          int64_t total_in_bytes, uint8_t *d_out, const int64_t *d_out_off,
                          int64_t *d_out_len, int64_t *d_status, void *d_scratch,
                          size_t scratch_bytes, cudaStream_t stream, int sm_count) {
    if (n_units <= 0) return cudaSuccess;
    ScratchLayout L = scratch_layout(n_units, total_in_bytes, dec_tile_in());
    if (scratch_bytes < L.total) return cudaErrorInvalidValue;
    cudaError_t e = plan(d_in_off, d_out_off, n_units, dec_tile_in(), dec_small_max(), 1, L, d_scratch, d_out_len,
                         d_status, stream);
    if (e != cudaSuccess) return e;
    return launch_b64dec_run(d_in, d_in_off, n_units, total_in_bytes, d_out, d_out_off, d_out_len,
                             d_status, d_scratch, scratch_bytes, stream, sm_count);
}

// Decode kernels over a tile table that is already in d_scratch (built by the plan kernels
// or by plan_tiles_host); d_out_len / d_status must already hold their initial values.
cudaError_t launch_b64dec_run(const char *d_in, const int64_t *d_in_off, int64_t n_units,
                              int64_t total_in_bytes, uint8_t *d_out, const int64_t *d_out_off,
                              int64_t *d_out_len, int64_t *d_status, void *d_scratch,
                              size_t scratch_bytes, cudaStream_t stream, int sm_count) {
    if (n_units <= 0) return cudaSuccess;
    ScratchLayout L = scratch_layout(n_units, total_in_bytes, dec_tile_in());
    if (scratch_bytes < L.total) return cudaErrorInvalidValue;
    char *base = static_cast<char *>(d_scratch);
    const DecVariant &dv = kDec[dec_variant()];
    int occ = 1;
    cudaError_t e = kernel_ready(reinterpret_cast<const void *>(dv.fn), dv.threads(), dv.dyn(), &occ);
    if (e != cudaSuccess) return e;
    int64_t grid = (int64_t)sm_count * occ;
    if (grid > (L.max_tiles + dv.groups - 1) / dv.groups) grid = (L.max_tiles + dv.groups - 1) / dv.groups;
    if (grid < 1) grid = 1;
    e = launch_after_kernel(dv.fn, dim3((unsigned)grid), dim3(dv.threads()), dv.dyn(), stream, d_in, d_out,
                            reinterpret_cast<const TileDesc *>(base + L.desc_off), reinterpret_cast<int *>(base),
                            d_in_off, n_units, d_out_len, d_status);
    if (e != cudaSuccess) return e;
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    int64_t sg = (n_units + 127) / 128;
    if (sg > (int64_t)sm_count * 16) sg = (int64_t)sm_count * 16;
    b64dec_slow<<<(unsigned)sg, 128, 0, stream>>>(d_in, d_in_off, n_units, d_out, d_out_off,
                                                  d_out_len, d_status);
    return cudaGetLastError();
}

}  // namespace kvedge
