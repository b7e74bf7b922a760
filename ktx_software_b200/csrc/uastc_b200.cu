// uastc_b200.cu -- CUDA kernels (sm_100a) and the C ABI declared in include/ktx_b200_uastc.h.
//
// Transcode (reference hot loop #3, XC/basisu_transcoder.cpp:10128-10305): one thread per 16-byte block,
// 128-bit coalesced loads/stores, grid sized in whole waves of the SM count, grid-stride over the span.
// Host variant: the spans are cut into slabs that stream H2D -> kernel -> D2H through a ring of device
// buffers on three streams so PCIe in, compute and PCIe out overlap.
#include <cuda_runtime.h>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>
#include <algorithm>
#include <stdio.h>

#include "../../include/ktx_b200_uastc.h"
#include "uastc_to_bc7.cuh"
#include "uastc_to_astc.cuh"
#include "uastc_encode.cuh"
#include "uastc_rdo.cuh"
#include "uastc_to_rgba.cuh"
#include "uastc_to_bcn.cuh"

namespace {

thread_local std::string g_err;
std::atomic<uint64_t> g_launches{0};

int fail(int code, const char* what, cudaError_t e = cudaSuccess) {
  g_err = what;
  if (e != cudaSuccess) { g_err += ": "; g_err += cudaGetErrorString(e); }
  return code;
}
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(e_ == cudaErrorMemoryAllocation ? B200_E_OUT_OF_MEMORY : B200_E_CUDA, #call, e_); } while (0)

struct DeviceCtx {
  int device = -1;
  int sm_count = 0;
  bool ready = false;
};
std::mutex g_mu;
DeviceCtx g_ctx;

// ---------------------------------------------------------------------------------------------------------
// kernels

template <int TARGET>  // 0 = BC7, 1 = ASTC
__global__ void __launch_bounds__(256) k_transcode(const uint4* __restrict__ src, uint4* __restrict__ dst, uint64_t n_blocks,
                                                   unsigned long long* __restrict__ first_bad) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_blocks; i += stride) {
    const uint4 v = __ldg(src + i);
    ub::Bits128 s, d;
    s.lo = (uint64_t)v.x | ((uint64_t)v.y << 32);
    s.hi = (uint64_t)v.z | ((uint64_t)v.w << 32);
    const bool ok = TARGET ? ub::transcode_block_to_astc(s, d) : ub::transcode_block_to_bc7(s, d);
    uint4 o;
    o.x = (uint32_t)d.lo; o.y = (uint32_t)(d.lo >> 32); o.z = (uint32_t)d.hi; o.w = (uint32_t)(d.hi >> 32);
    dst[i] = o;
    if (!ok && first_bad) atomicMin(first_bad, (unsigned long long)i);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Tiled transcode: the streaming kernel.  A CTA owns tiles of XT_TILE consecutive blocks (16 KB in, 16 KB out).
//   1. one thread issues a TMA bulk copy (cp.async.bulk, mbarrier completion) of the NEXT tile while the current
//      one is processed (double-buffered);
//   2. the tile is binned by UASTC mode in shared memory (warp-aggregated counting sort of block indices), so that
//      a warp works on blocks of ONE mode and runs that mode's fully specialised unpack/pack code -- every bit
//      offset of a UASTC mode is a compile-time constant -- instead of diverging over the 19-way switch;
//   3. each result overwrites its block's slot of the shared-memory tile, which leaves as one TMA bulk store.
// Global traffic is exactly the algorithmic 32 B per block, fully coalesced by construction.
#ifndef XT_THREADS
#define XT_THREADS 1024
#endif
#ifndef XT_TILE
#define XT_TILE 4096
#endif
#ifndef XT_SMEM_TABLES
#define XT_SMEM_TABLES true   // endpoint-unquantisation and p-bit tables are read from a per-CTA shared-memory copy
#endif
#ifndef XT_DYNAMIC
#define XT_DYNAMIC 1   // hint-driven targets: warps claim 32-block runs of the sorted tile dynamically (0: fixed share per warp)
#endif
#ifndef XT_CTAS_PER_SM
#define XT_CTAS_PER_SM 1
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <int TARGET>
__device__ __forceinline__ bool transcode_dispatch(uint32_t mode, const ub::Bits128& s, ub::Bits128& d) {
  switch (mode) {
#define XC_CASE(M) case M: return ub::transcode_known_mode<TARGET, M, XT_SMEM_TABLES>(s, d);
    XC_CASE(0) XC_CASE(1) XC_CASE(2) XC_CASE(3) XC_CASE(4) XC_CASE(5) XC_CASE(6) XC_CASE(7) XC_CASE(8) XC_CASE(9)
    XC_CASE(10) XC_CASE(11) XC_CASE(12) XC_CASE(13) XC_CASE(14) XC_CASE(15) XC_CASE(16) XC_CASE(17) XC_CASE(18)
#undef XC_CASE
    default: d.lo = 0; d.hi = 0; return false;
  }
}

__device__ __forceinline__ bool texel_form_dispatch(uint32_t mode, const ub::Bits128& s, ub::TexelForm& f) {
  switch (mode) {
#define XC_CASE(M) case M: return ub::texel_form_t<M, XT_SMEM_TABLES>(s, f);
    XC_CASE(0) XC_CASE(1) XC_CASE(2) XC_CASE(3) XC_CASE(4) XC_CASE(5) XC_CASE(6) XC_CASE(7) XC_CASE(8) XC_CASE(9)
    XC_CASE(10) XC_CASE(11) XC_CASE(12) XC_CASE(13) XC_CASE(14) XC_CASE(15) XC_CASE(16) XC_CASE(17) XC_CASE(18)
#undef XC_CASE
    default: ub::texel_form_clear(f); return false;   // reserved mode code: the form decodes to zeros
  }
}

// Uncompressed targets (TARGET 2): the blocks of ONE image in raster order are written as texels into a w x h image
// (the cRGBA32 / cRGB565 / cBGR565 / cRGBA4444 cases of transcode_slice, XC:10208-10292: rows and columns past the
// image's own size are clipped).
struct XtImage {
  uint32_t w, h;         // texels
  uint32_t bx;           // blocks per row
  uint32_t pitch_px;     // destination row pitch in texels
  uint32_t kind;         // 13 = RGBA32, 14 = RGB565, 15 = BGR565, 16 = RGBA4444 (ktx_transcode_fmt_e values)
  uint32_t vec_ok;       // destination base and pitch allow 16-byte (RGBA32) / 8-byte (16-bit targets) row stores
  // TARGET 3 (hint-driven block targets) uses only: kind = 0 ETC1_RGB, 1 ETC2_RGBA, 2 BC1_RGB, 3 BC3_RGBA, 4 BC4_R, 5 BC5_RG
  uint32_t hq;           // cDecodeFlagsHighQuality
  uint32_t chan0, chan1; // channels of BC4 / BC5 (transcode_image, XC:10417-10438)
};

#ifndef XT_DECODE_STAGED
#define XT_DECODE_STAGED 1   // 1: texel rows leave through a shared-memory staging tile (coalesced); 0: straight from registers
#endif
__host__ __device__ constexpr int xt_tile(int target) { return (target == 2 && XT_DECODE_STAGED) ? 2048 : XT_TILE; }   // staging holds 64 B per block
__host__ __device__ constexpr int xt_data_tiles(int target) { return (target == 2 && XT_DECODE_STAGED) ? 6 : 2; }
// TARGET 3 adds a TILE x 8-byte output tile for the 8-byte targets (their results cannot overwrite the 16-byte input slots)
__host__ __device__ constexpr size_t xt_data_bytes(int target) { return (size_t)xt_data_tiles(target) * xt_tile(target) * 16 + (target == 3 ? (size_t)xt_tile(target) * 8 : 0); }
constexpr size_t xt_smem_bytes(int target) { return xt_data_bytes(target) + 2 * 8 + 72 * 4 + (size_t)xt_tile(target) * 2; }

__device__ __forceinline__ void store_block_texels(uint8_t* __restrict__ dst, const XtImage& im, uint64_t block_index, const uint32_t* px) {
  const uint32_t by = (uint32_t)(block_index / im.bx), bxi = (uint32_t)(block_index - (uint64_t)by * im.bx);
  const uint32_t x0 = bxi * 4, y0 = by * 4;
  const bool fast = im.vec_ok && x0 + 4 <= im.w && y0 + 4 <= im.h;
  const uint32_t bpp = im.kind == 13 ? 4u : 2u;
  uint8_t* row = dst + ((size_t)y0 * im.pitch_px + x0) * bpp;
#pragma unroll
  for (int y = 0; y < 4; y++) {
    uint32_t q[4];
#pragma unroll
    for (int x = 0; x < 4; x++) q[x] = im.kind == 13 ? px[4 * y + x] : ub::pack_texel16(px[4 * y + x], im.kind);
    if (fast) {
      if (im.kind == 13) *(uint4*)row = make_uint4(q[0], q[1], q[2], q[3]);
      else *(uint2*)row = make_uint2(q[0] | (q[1] << 16), q[2] | (q[3] << 16));
    } else if (y0 + y < im.h) {
#pragma unroll
      for (int x = 0; x < 4; x++)
        if (x0 + x < im.w) { if (im.kind == 13) ((uint32_t*)row)[x] = q[x]; else ((uint16_t*)row)[x] = (uint16_t)q[x]; }
    }
    row += (size_t)im.pitch_px * bpp;
  }
}

// Row pieces of block `slot` of the staging tile -> the image.  Full blocks of an aligned destination take one
// 16-byte (RGBA32) or 8-byte (16-bit targets) store per texel row; edge blocks and unaligned images go texel by texel.
__device__ __forceinline__ void store_block_rows(uint8_t* __restrict__ dst, const XtImage& im, uint64_t block_index, const uint4* s_out,
                                                 uint32_t tile, uint32_t slot) {
  const uint32_t by = (uint32_t)(block_index / im.bx), bxi = (uint32_t)(block_index - (uint64_t)by * im.bx);
  const uint32_t x0 = bxi * 4, y0 = by * 4;
  const bool fast = im.vec_ok && x0 + 4 <= im.w && y0 + 4 <= im.h;
  if (im.kind == 13) {
    uint8_t* row = dst + ((size_t)y0 * im.pitch_px + x0) * 4;
#pragma unroll
    for (int y = 0; y < 4; y++) {
      const uint4 v = s_out[y * tile + slot];
      if (fast) *(uint4*)row = v;
      else if (y0 + y < im.h) {
        const uint32_t q[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int x = 0; x < 4; x++) if (x0 + x < im.w) ((uint32_t*)row)[x] = q[x];
      }
      row += (size_t)im.pitch_px * 4;
    }
  } else {
    uint8_t* row = dst + ((size_t)y0 * im.pitch_px + x0) * 2;
#pragma unroll
    for (int y = 0; y < 4; y++) {
      const uint2 v = ((const uint2*)s_out)[y * tile + slot];
      if (fast) *(uint2*)row = v;
      else if (y0 + y < im.h) {
        const uint32_t q[4] = {v.x & 0xFFFFu, v.x >> 16, v.y & 0xFFFFu, v.y >> 16};
#pragma unroll
        for (int x = 0; x < 4; x++) if (x0 + x < im.w) ((uint16_t*)row)[x] = (uint16_t)q[x];
      }
      row += (size_t)im.pitch_px * 2;
    }
  }
}

__device__ __forceinline__ bool hinted_form_dispatch(uint32_t mode, const ub::Bits128& s, ub::HintedForm& h) {
  switch (mode) {
#define XC_CASE(M) case M: return ub::hinted_form_t<M, XT_SMEM_TABLES>(s, h);
    XC_CASE(0) XC_CASE(1) XC_CASE(2) XC_CASE(3) XC_CASE(4) XC_CASE(5) XC_CASE(6) XC_CASE(7) XC_CASE(8) XC_CASE(9)
    XC_CASE(10) XC_CASE(11) XC_CASE(12) XC_CASE(13) XC_CASE(14) XC_CASE(15) XC_CASE(16) XC_CASE(17) XC_CASE(18)
#undef XC_CASE
    default: ub::texel_form_clear(h.f); h.mode = ub::UASTC_MODE_SOLID; h.hints = 0; h.bc1_sels = 0; h.solid = 0; return false;
  }
}

// 0 = BC7, 1 = ASTC (dst = blocks, written through the shared-memory tile), 2 = texels (dst = image, im),
// 3 = the hint-driven block targets (KIND = ktx_transcode_fmt_e 0..5: ETC1 / ETC2 / BC1 / BC3 / BC4 / BC5, 6 / 7 = ETC2 EAC
// R11 / RG11; 8 or 16 bytes per block).  KIND is a template parameter so each target's kernel holds only its own packer (one kernel with a run-time
// switch over six packers paid the register and spill cost of the largest, encode_bc1, in every target).
template <int TARGET, int KIND = 0>
__global__ void __launch_bounds__(XT_THREADS) k_transcode_tiled(const uint4* __restrict__ src, uint4* __restrict__ dst, uint64_t n_blocks,
                                                                unsigned long long* __restrict__ first_bad, const XtImage im) {
  constexpr int TILE = xt_tile(TARGET), PER_THREAD = TILE / XT_THREADS;
  extern __shared__ __align__(128) unsigned char xt_smem[];
  uint4 (*s_in)[TILE] = (uint4 (*)[TILE])xt_smem;  // [2][TILE]; a block's result overwrites its own input slot
  uint4* s_out = (uint4*)(xt_smem + 2 * (size_t)TILE * 16);   // TARGET 2 only: [4 texel rows][TILE] row pieces of 16 (or 8) bytes
  uint2* s_out8 = (uint2*)(xt_smem + 2 * (size_t)TILE * 16);  // TARGET 3, 8-byte targets: the output tile
  uint64_t* s_bar = (uint64_t*)(xt_smem + xt_data_bytes(TARGET));
  constexpr bool out8 = TARGET == 3 && (KIND == 0 || KIND == 2 || KIND == 4 || KIND == 6);
  uint32_t* s_hist = (uint32_t*)(s_bar + 2);
  // BC1 / BC3: the cost of a block is set by which BC1 path its hint bits select (hint0: endpoints + weights as they are;
  // hint1: one least-squares fit on the block's weights; none: the full encode_bc1), so the bins are (path, mode) pairs --
  // 3 x 19 of them, scanned as two bins per lane -- and a warp runs one path of one mode.
  constexpr bool WIDE = TARGET == 3 && (KIND == 2 || KIND == 3);
  constexpr uint32_t NBINS = WIDE ? 64 : 32;
  // Dynamic claiming of the sorted tile (see the transcode loop) for the hint-driven targets only.  Same-box A/B on BC7
  // (tools/ab_xcode.py, bench.py --workload c5; profiles/r3d_ab_claiming.txt): +5..12 % on image-ordered input (tiles whose
  // bins differ in size: C2's blocks, the golden corpus) but -6 % on C5's i.i.d. block mix, where every round of a fixed
  // share is homogeneous already; a variant that claims only the last of four rounds loses the same 6 % there, so the cost
  // is the claim loop's code, not the schedule.  BC7 / ASTC / texels keep the fixed share until that is understood.
  constexpr bool DYN = XT_DYNAMIC && TARGET == 3;
  uint32_t* s_next = s_hist + NBINS;                    // next unclaimed position of the sorted tile (see the transcode loop)
  uint16_t* s_order = (uint16_t*)(s_hist + NBINS + 2);

  const uint32_t tid = threadIdx.x, lane = tid & 31u;
  const uint64_t n_tiles = (n_blocks + TILE - 1) / TILE;
  uint64_t tile = blockIdx.x;
  if (tile >= n_tiles) return;
  auto tile_count = [&](uint64_t t) { return (uint32_t)min((uint64_t)TILE, n_blocks - t * TILE); };

  if (tid < NBINS) s_hist[tid] = 0;
  if (XT_SMEM_TABLES) ub::shared_tables_load(tid, XT_THREADS);
  if (tid == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
    const uint32_t bytes = tile_count(tile) * 16u;
    mbar_expect_tx(&s_bar[0], bytes);
    bulk_g2s(s_in[0], src + tile * TILE, bytes, &s_bar[0]);
  }
  __syncthreads();

  uint32_t buf = 0, phase0 = 0, phase1 = 0;
  for (; tile < n_tiles; tile += gridDim.x) {
    const uint32_t cnt = tile_count(tile);
    const uint64_t next = tile + gridDim.x;
    if (tid == 0 && (TARGET == 3 || next < n_tiles)) {  // prefetch the next tile into the other buffer once its bulk store has drained it
      bulk_wait_read0();   // (TARGET 3: also frees the single 8-byte output tile before this iteration refills it)
      if (next < n_tiles) {
        const uint32_t bytes = tile_count(next) * 16u;
        mbar_expect_tx(&s_bar[buf ^ 1], bytes);
        bulk_g2s(s_in[buf ^ 1], src + next * TILE, bytes, &s_bar[buf ^ 1]);
      }
    }
    if (DYN && tid == 32) *s_next = 0;   // (every warp left the previous tile's loop at the barrier that ended it)
    mbar_wait(&s_bar[buf], buf ? phase1 : phase0);   // every thread waits for the tile itself: no CTA barrier needed here
    if (buf) phase1 ^= 1; else phase0 ^= 1;

    // ---- bin by mode: rank inside the bin from a warp-aggregated shared atomic (s_hist is zero on entry)
    uint32_t my_mode[PER_THREAD], my_rank[PER_THREAD];
#pragma unroll
    for (int k = 0; k < PER_THREAD; k++) {
      const uint32_t i = tid + k * XT_THREADS;
      const bool live = i < cnt;
      const uint32_t w0 = s_in[buf][live ? i : 0].x;
      uint32_t m = live ? (uint32_t)ub::UBT(uastc_huff_modes)[w0 & 127u] : 31u;
      if (WIDE) {   // m becomes the (path, mode) bin: 0..56, 62 = invalid mode code, 63 = no block
        if (!live) m = 63u;
        else if (m >= ub::UASTC_TOTAL_MODES) m = 62u;
        else {
          const uint32_t len = ub::md_huff_len(ub::UBT(uastc_mode_desc)[m]);
          const uint32_t has0 = ub::UBT(uastc_mode_has_bc1_hint0)[m], has1 = ub::UBT(uastc_mode_has_bc1_hint1)[m];
          const uint32_t h0 = has0 ? ((w0 >> len) & 1u) : 0u, h1 = has1 ? ((w0 >> (len + has0)) & 1u) : 0u;
          m += 19u * ((!im.hq && h0) ? 0u : (h1 ? 1u : 2u));
        }
      }
      const uint32_t peers = __match_any_sync(0xFFFFFFFFu, m);
      const uint32_t leader = __ffs(peers) - 1;
      uint32_t base = 0;
      if (lane == leader) base = atomicAdd(&s_hist[m], __popc(peers));
      base = __shfl_sync(0xFFFFFFFFu, base, leader);
      my_mode[k] = m;
      my_rank[k] = base + __popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();
    // every warp scans the 32 bin counts for itself (lane = bin): no second barrier, no shared-memory round trip
    uint32_t bin_start, bin_start_hi = 0;
    {
      const uint32_t c = s_hist[lane];
      uint32_t v = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, v, o); if ((int)lane >= o) v += t; }
      bin_start = v - c;
      if (WIDE) {   // bins 32..63 follow
        const uint32_t total_lo = __shfl_sync(0xFFFFFFFFu, v, 31);
        const uint32_t c1 = s_hist[32 + lane];
        uint32_t v1 = c1;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, v1, o); if ((int)lane >= o) v1 += t; }
        bin_start_hi = total_lo + v1 - c1;
      }
    }
#pragma unroll
    for (int k = 0; k < PER_THREAD; k++) {
      const uint32_t i = tid + k * XT_THREADS;
      uint32_t start = __shfl_sync(0xFFFFFFFFu, bin_start, WIDE ? (my_mode[k] & 31u) : my_mode[k]);
      if (WIDE) { const uint32_t hi = __shfl_sync(0xFFFFFFFFu, bin_start_hi, my_mode[k] & 31u); if (my_mode[k] >= 32u) start = hi; }
      if (i < cnt) s_order[start + my_rank[k]] = (uint16_t)i;
    }
    __syncthreads();
    if (tid < NBINS) s_hist[tid] = 0;   // for the next tile; ordered before its atomics by the barrier that ends this one

    // ---- transcode in bin order: a warp sees one mode (two at a bin boundary)
    // Warps claim runs of 32 sorted positions from a shared counter instead of a fixed share: modes differ several-fold in
    // cost, and with a fixed share every warp waits at the tile's last barrier for the one that drew the expensive bins.
#pragma unroll 1
    for (int k = 0; DYN || k < PER_THREAD; k++) {
      uint32_t j = tid + k * XT_THREADS;
      if (DYN) {
        uint32_t c = 0;
        if (lane == 0) c = atomicAdd(s_next, 32u);
        c = __shfl_sync(0xFFFFFFFFu, c, 0);
        if (c >= cnt) break;
        j = c + lane;
      }
      if (j < cnt) {
        const uint32_t idx = s_order[j];
        const uint4 v = s_in[buf][idx];
        const uint32_t m = (uint32_t)ub::UBT(uastc_huff_modes)[v.x & 127u];
        ub::Bits128 s, d;
        s.lo = (uint64_t)v.x | ((uint64_t)v.y << 32);
        s.hi = (uint64_t)v.z | ((uint64_t)v.w << 32);
        bool ok;
        if (TARGET == 2) {
          uint32_t px[16];
          ub::TexelForm f;
          ok = texel_form_dispatch(m, s, f);   // mode-specialised
          ub::texels_from_form(f, px);          // shared by all modes: lanes of mixed warps reconverge here
          // texel rows of the block go to the staging tile in BLOCK order (not sorted order), one piece per row
          if (!XT_DECODE_STAGED) store_block_texels((uint8_t*)dst, im, tile * TILE + idx, px);
          else if (im.kind == 13) {
#pragma unroll
            for (int y = 0; y < 4; y++) s_out[y * TILE + idx] = make_uint4(px[4 * y], px[4 * y + 1], px[4 * y + 2], px[4 * y + 3]);
          } else {
#pragma unroll
            for (int y = 0; y < 4; y++) {
              uint32_t q[4];
#pragma unroll
              for (int x = 0; x < 4; x++) q[x] = ub::pack_texel16(px[4 * y + x], im.kind);
              ((uint2*)s_out)[y * TILE + idx] = make_uint2(q[0] | (q[1] << 16), q[2] | (q[3] << 16));
            }
          }
        } else if (TARGET == 3) {
          ub::HintedForm h;
          ok = hinted_form_dispatch(m, s, h);   // mode-specialised
          if (ok) ub::hinted_target_from_form(h, (uint32_t)KIND, im.hq != 0, im.chan0, im.chan1, d);   // shared by all modes
          else { d.lo = 0; d.hi = 0; }
          if (out8) s_out8[idx] = make_uint2((uint32_t)d.lo, (uint32_t)(d.lo >> 32));
          else s_in[buf][idx] = make_uint4((uint32_t)d.lo, (uint32_t)(d.lo >> 32), (uint32_t)d.hi, (uint32_t)(d.hi >> 32));
        } else {
          ok = transcode_dispatch<TARGET>(m, s, d);
          uint4 o;
          o.x = (uint32_t)d.lo; o.y = (uint32_t)(d.lo >> 32); o.z = (uint32_t)d.hi; o.w = (uint32_t)(d.hi >> 32);
          s_in[buf][idx] = o;
        }
        if (!ok && first_bad) atomicMin(first_bad, (unsigned long long)(tile * TILE + idx));
      }
    }
    if (TARGET != 2) fence_proxy_async();  // make the generic-proxy writes of the tile visible to the bulk-copy engine
    __syncthreads();
    if (TARGET != 2 && !out8 && tid == 0) bulk_s2g(dst + tile * TILE, s_in[buf], cnt * 16u);
    if (out8 && tid == 0) {   // bulk copies move multiples of 16 bytes: an odd last block leaves by a plain store
      uint2* o8 = (uint2*)dst + tile * TILE;
      if (cnt & ~1u) bulk_s2g(o8, s_out8, (cnt & ~1u) * 8u);
      if (cnt & 1u) o8[cnt - 1] = s_out8[cnt - 1];
    }
    if (TARGET == 2 && XT_DECODE_STAGED) {
      // consecutive threads take consecutive blocks of the image: their row pieces are consecutive in global memory
      for (uint32_t j = tid; j < cnt; j += XT_THREADS) store_block_rows((uint8_t*)dst, im, tile * TILE + j, s_out, TILE, j);
      // (the barrier after the next tile's binning orders these reads before the next writes of s_out)
    }
    buf ^= 1;
  }
  if (tid == 0) bulk_wait0();
}

// Encode (reference hot loop #1, ENC/basisu_comp.cpp:1519-1562 -> encode_uastc): one thread per 4x4 block, so a warp
// is a group of 32 neighbouring blocks that walk the mode list in lock-step.  The 4x4 gather with edge replication
// (image::extract_block_clamped, ENC/basisu_enc.h:3081) is fused into the load.  MAXC bounds the candidate list:
// 21 at levels 0-3 (partition estimators on), 170 at level 4.
#ifndef ENC_THREADS
#define ENC_THREADS 1024  // threads per encode CTA: one CTA per SM, 32 warps, 64 registers per thread
#endif
#ifndef ENC_FINE_SYNC
#define ENC_FINE_SYNC true  // barriers between the per-mode stages inside candidate generation, not only between the phases
#endif
#ifndef ENC_FINE_SYNC
#define ENC_FINE_SYNC true  // barriers between the per-mode stages inside candidate generation, not only between the phases
#endif
#ifndef ENC_MIN_CTAS
#define ENC_MIN_CTAS 1    // resident encode CTAs per SM (register cap = 65536 / (ENC_THREADS * ENC_MIN_CTAS))
#endif
// The warps of a CTA walk the three phases of encode_uastc together (barrier between phases): all warps of the CTA
// -- with one CTA per SM, all warps of the SM -- then execute the same few tens of KB of code at a time, which is
// what the 32 KB instruction cache can hold; free-running warps spread over the whole ~300 KB kernel.
// Thread index -> block of an image.  Consecutive groups of 32 threads (a warp) take a 4 x 8 patch of blocks (16 x 32
// texels) instead of a 32 x 1 strip, so the lanes of a warp see similar content and diverge less; and consecutive
// groups of 1024 threads (a full-size CTA) take a 32 x 32 tile of blocks, so the warps that meet at the stage barriers
// carry similar work.  A permutation of [0, bx * by): the 32 x 32 tiles first, then the columns right of the last
// whole tile, then the rows below the last whole band, both in 8 x 4 patches as far as they go.
__device__ __forceinline__ uint64_t patch_index(uint64_t i, uint32_t bx, uint32_t by) {   // 8 x 4 patches inside a bx x by rectangle
  const uint64_t band = 4ull * bx;
  if (i >= (uint64_t)(by / 4) * band) return i;
  const uint32_t r = (uint32_t)(i / band), j = (uint32_t)(i % band), bx8 = bx & ~7u;
  uint32_t x, y;
  if (j < 4 * bx8) { x = (j >> 5) * 8 + (j & 7u); y = (j & 31u) >> 3; }
  else { const uint32_t jj = j - 4 * bx8, rc = bx - bx8; x = bx8 + jj % rc; y = jj / rc; }
  return (uint64_t)(r * 4 + y) * bx + x;
}
#ifndef ENC_TILE_W
#define ENC_TILE_W 32   // blocks per CTA tile row (tile = 1024 blocks)
#endif
#ifndef ENC_PATCH_W
#define ENC_PATCH_W 4   // blocks per warp patch row (patch = 32 blocks: 4 x 8, i.e. 16 x 32 texels)
#endif
__device__ __forceinline__ uint64_t block_of_thread(uint64_t i, uint32_t bx, uint32_t by) {
  constexpr uint32_t TW = ENC_TILE_W, TH = 1024 / TW, PW = ENC_PATCH_W, PH = 32 / PW, PPR = TW / PW;
  const uint32_t tpb = bx / TW, bands = by / TH;            // whole tiles per band, whole bands
  const uint64_t n_tiled = (uint64_t)tpb * bands * 1024;
  if (i < n_tiled) {
    const uint32_t tile = (uint32_t)(i >> 10), within = (uint32_t)(i & 1023u), warp = within >> 5, lane = within & 31u;
    const uint32_t x = (tile % tpb) * TW + (warp % PPR) * PW + (lane % PW), y = (tile / tpb) * TH + (warp / PPR) * PH + (lane / PW);
    return (uint64_t)y * bx + x;
  }
  uint64_t q = i - n_tiled;
  const uint32_t rc = bx - tpb * TW, rows_t = bands * TH;   // ragged columns beside the tiles
  if (q < (uint64_t)rc * rows_t) {
    const uint64_t k = patch_index(q, rc, rows_t);
    return (k / rc) * bx + tpb * TW + (k % rc);
  }
  q -= (uint64_t)rc * rows_t;
  return (uint64_t)rows_t * bx + patch_index(q, bx, by - rows_t);   // rows below the last whole band
}

template <int MAXC, typename LOAD>
__device__ __forceinline__ void encode_phased(uint64_t n, uint32_t flags, uint4* __restrict__ out, LOAD load_px, uint32_t bx = 0, uint32_t by = 0) {
  ub::Candidate cands[MAXC];
  uint64_t overall[MAXC];
  float uerr[MAXC];
  for (uint64_t base = (uint64_t)blockIdx.x * blockDim.x; base < n; base += (uint64_t)gridDim.x * blockDim.x) {  // uniform per CTA
    const uint64_t t = base + threadIdx.x;
    const bool live = t < n;
    const uint64_t i = (live && bx) ? block_of_thread(t, bx, by) : t;   // bx = 0: blocks are not laid out as an image
    uint32_t px[16];
    ub::EncState st;
    if (live) load_px(i, px);
    ub::enc_phase_candidates(px, flags, cands, st, live, ENC_FINE_SYNC);
    __syncthreads();
    ub::enc_phase_select(px, flags, cands, overall, uerr, st, live);
    __syncthreads();
    const ub::Bits128 b = ub::enc_phase_finish(px, flags, cands, st, live, ENC_FINE_SYNC);
    if (live) out[i] = make_uint4((uint32_t)b.lo, (uint32_t)(b.lo >> 32), (uint32_t)b.hi, (uint32_t)(b.hi >> 32));
    __syncthreads();
  }
}

template <int MAXC>
__global__ void __launch_bounds__(ENC_THREADS, ENC_MIN_CTAS) k_encode_image(const uint8_t* __restrict__ img, uint32_t w, uint32_t h, uint32_t pitch,
                                                      uint4* __restrict__ out, uint32_t bx, uint32_t by, uint32_t flags,
                                                      unsigned int* __restrict__ any_alpha) {
  encode_phased<MAXC>((uint64_t)bx * by, flags, out, [&](uint64_t i, uint32_t* px) {
    const uint32_t x0 = (uint32_t)(i % bx) * 4, y0 = (uint32_t)(i / bx) * 4;
    if (x0 + 4 <= w && y0 + 4 <= h && !(pitch & 15u)) {
#pragma unroll
      for (uint32_t y = 0; y < 4; y++) {
        const uint4 v = __ldg((const uint4*)(img + (size_t)(y0 + y) * pitch + (size_t)x0 * 4));
        px[4 * y] = v.x; px[4 * y + 1] = v.y; px[4 * y + 2] = v.z; px[4 * y + 3] = v.w;
      }
    } else {
      for (uint32_t y = 0; y < 4; y++)
        for (uint32_t x = 0; x < 4; x++) {
          const uint32_t sx = min(x0 + x, w - 1), sy = min(y0 + y, h - 1);
          const uint8_t* q = img + (size_t)sy * pitch + (size_t)sx * 4;
          px[4 * y + x] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
        }
    }
    uint32_t amin = 255;
#pragma unroll
    for (uint32_t k = 0; k < 16; k++) amin = min(amin, px[k] >> 24);
    if (amin < 255 && any_alpha) atomicOr(any_alpha, 1u);
  }, bx, by);
}

// Several images in one launch (a mip chain, the faces of a cube map): block i belongs to the image whose
// [first_block, first_block + blocks) range holds it.
struct EncImageDesc {
  const uint8_t* img;
  uint4* out;
  uint32_t w, h, pitch, bx;
  uint64_t first_block;
};

template <int MAXC>
__global__ void __launch_bounds__(ENC_THREADS, ENC_MIN_CTAS) k_encode_images(const EncImageDesc* __restrict__ descs, uint32_t n_images, uint64_t total_blocks,
                                                                             uint32_t flags, unsigned int* __restrict__ any_alpha) {
  ub::Candidate cands[MAXC];
  uint64_t overall[MAXC];
  float uerr[MAXC];
  for (uint64_t base = (uint64_t)blockIdx.x * blockDim.x; base < total_blocks; base += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t i = base + threadIdx.x;
    const bool live = i < total_blocks;
    uint32_t px[16];
    ub::EncState st;
    uint4* dst = nullptr;
    if (live) {
      uint32_t lo = 0, hi = n_images - 1;  // last image whose first_block <= i
      while (lo < hi) { const uint32_t mid = (lo + hi + 1) >> 1; if (descs[mid].first_block <= i) lo = mid; else hi = mid - 1; }
      const EncImageDesc d = descs[lo];
      const uint64_t k = block_of_thread(i - d.first_block, d.bx, (d.h + 3) / 4);
      dst = d.out + k;
      const uint32_t x0 = (uint32_t)(k % d.bx) * 4, y0 = (uint32_t)(k / d.bx) * 4;
      uint32_t amin = 255;
      for (uint32_t y = 0; y < 4; y++)
        for (uint32_t x = 0; x < 4; x++) {
          const uint32_t sx = min(x0 + x, d.w - 1), sy = min(y0 + y, d.h - 1);
          const uint32_t v = __ldg((const uint32_t*)(d.img + (size_t)sy * d.pitch + (size_t)sx * 4));
          px[4 * y + x] = v;
          amin = min(amin, v >> 24);
        }
      if (amin < 255 && any_alpha) atomicOr(any_alpha, 1u);
    }
    ub::enc_phase_candidates(px, flags, cands, st, live, ENC_FINE_SYNC);
    __syncthreads();
    ub::enc_phase_select(px, flags, cands, overall, uerr, st, live);
    __syncthreads();
    const ub::Bits128 b = ub::enc_phase_finish(px, flags, cands, st, live, ENC_FINE_SYNC);
    if (live) *dst = make_uint4((uint32_t)b.lo, (uint32_t)(b.lo >> 32), (uint32_t)b.hi, (uint32_t)(b.hi >> 32));
    __syncthreads();
  }
}

template <int MAXC>
__global__ void __launch_bounds__(ENC_THREADS, ENC_MIN_CTAS) k_encode_blocks(const uint4* __restrict__ src, uint4* __restrict__ out, uint64_t n, uint32_t flags) {
  encode_phased<MAXC>(n, flags, out, [&](uint64_t i, uint32_t* px) {
#pragma unroll
    for (uint32_t y = 0; y < 4; y++) {
      const uint4 v = __ldg(src + i * 4 + y);
      px[4 * y] = v.x; px[4 * y + 1] = v.y; px[4 * y + 2] = v.z; px[4 * y + 3] = v.w;
    }
  });
}

// ---------------------------------------------------------------------------------------------------------
// RDO (reference hot loop #2, ENC/basisu_uastc_enc.cpp:3825-4165).  Three kernels:
//   k_rdo_prep    one warp per block, lane = weight level: the per-texel error tables of uastc_rdo.cuh (parallel);
//   k_rdo_chain   one CTA per job range: the serial walk.  256 threads score the (up to) 256 window candidates of
//                 the current block from its table in shared memory, a 64-bit shared atomicMin picks the winner
//                 (ties -> nearest earlier block, the reference's scan order), thread 0 commits it;
//   k_rdo_rehint  one thread per modified block: recompute the transcoding hints (parallel).
// The walk is cut into chunks of RDO_CHUNK steps so the tables of a chunk (2 KB per block) stay small.
struct RdoChain {
  uint32_t first, last;      // job range [first, last) in the concatenated block array
  uint64_t hist_off;         // first entry of this chain's selector history
  uint32_t hist_mask;
  uint32_t pad;
};
const int RDO_THREADS = 288;       // 256 trial threads + one warp that prepares the current block
const uint32_t RDO_CHUNK = 8192;   // steps per launch

__global__ void __launch_bounds__(256) k_gather_blocks(const uint8_t* __restrict__ img, uint32_t w, uint32_t h, uint32_t pitch,
                                                       uint4* __restrict__ out, uint32_t bx, uint32_t by, unsigned int* __restrict__ any_alpha) {
  const uint32_t n = bx * by;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t x0 = (i % bx) * 4, y0 = (i / bx) * 4;
    uint32_t amin = 255;
    for (uint32_t y = 0; y < 4; y++) {
      uint32_t r[4];
      for (uint32_t x = 0; x < 4; x++) {
        const uint32_t sx = min(x0 + x, w - 1), sy = min(y0 + y, h - 1);
        const uint8_t* q = img + (size_t)sy * pitch + (size_t)sx * 4;
        r[x] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
        amin = min(amin, r[x] >> 24);
      }
      out[(size_t)i * 4 + y] = make_uint4(r[0], r[1], r[2], r[3]);
    }
    if (amin < 255 && any_alpha) atomicOr(any_alpha, 1u);
  }
}

// Source texels of 1-4 components -> RGBA8 through libktx's component mapping (swizzle_to_rgba / copy_rgb_to_rgba,
// lib/src/basis_encode.cpp:100-144): output channel c takes source component sel_c - 1 (R..A = 1..4), 0x00 (5) or 0xFF (6).
// A streaming pass at HBM speed (<= 8 B per texel against the encoder's ~1.3 us of ALU work per block).
__global__ void __launch_bounds__(256) k_expand_swizzle(const uint8_t* __restrict__ src, uint32_t w, uint32_t rows, uint32_t src_pitch,
                                                        uint32_t comps, uint32_t sel, uint32_t* __restrict__ dst) {
  const uint64_t n = (uint64_t)w * rows;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t y = (uint32_t)(i / w), x = (uint32_t)(i - (uint64_t)y * w);
    const uint8_t* q = src + (size_t)y * src_pitch + (size_t)x * comps;
    uint32_t b[4] = {0, 0, 0, 0};
    for (uint32_t k = 0; k < comps; k++) b[k] = q[k];
    uint32_t o = 0;
#pragma unroll
    for (uint32_t c = 0; c < 4; c++) {
      const uint32_t sc = (sel >> (8 * c)) & 255u;
      const uint32_t v = sc <= 4 ? b[(sc - 1) & 3u] : (sc == 5 ? 0u : 255u);
      o |= v << (8 * c);
    }
    dst[i] = o;
  }
}

__device__ __forceinline__ ub::Bits128 to_bits(const uint4& v) {
  ub::Bits128 b;
  b.lo = (uint64_t)v.x | ((uint64_t)v.y << 32);
  b.hi = (uint64_t)v.z | ((uint64_t)v.w << 32);
  return b;
}
__device__ __forceinline__ uint4 from_bits(const ub::Bits128& b) {
  return make_uint4((uint32_t)b.lo, (uint32_t)(b.lo >> 32), (uint32_t)b.hi, (uint32_t)(b.hi >> 32));
}

// chunk-local table slot of block `bi` of chain `c`: c * RDO_CHUNK + (bi - first - step0)
__global__ void __launch_bounds__(256) k_rdo_prep(const uint4* __restrict__ blocks, const uint4* __restrict__ px, const RdoChain* __restrict__ chains,
                                                  uint32_t n_chains, uint32_t step0, ub::RdoBlockInfo* __restrict__ info,
                                                  ub::RdoPre* __restrict__ pre, uint32_t* __restrict__ tables,
                                                  unsigned int* __restrict__ err_flag, ub::RdoParams prm) {
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
  const uint32_t c = warp / RDO_CHUNK, k = warp % RDO_CHUNK;
  if (c >= n_chains) return;
  const uint32_t bi = chains[c].first + step0 + k;
  if (bi >= chains[c].last) return;
  uint32_t p[16];
#pragma unroll
  for (uint32_t y = 0; y < 4; y++) {
    const uint4 v = __ldg(px + (size_t)bi * 4 + y);
    p[4 * y] = v.x; p[4 * y + 1] = v.y; p[4 * y + 2] = v.z; p[4 * y + 3] = v.w;
  }
  const ub::Bits128 blk = to_bits(__ldg(blocks + bi));
  ub::RdoBlockInfo f;
  ub::UnpackedUastc u;
  const bool ok = ub::rdo_block_info(blk, f, u);
  if (lane == 0) { info[(size_t)c * RDO_CHUNK + k] = f; if (!ok) atomicOr(err_flag, 1u); }
  if (!ok || f.mode == ub::UASTC_MODE_SOLID) return;
  uint32_t* table = tables + ((size_t)c * RDO_CHUNK + k) * ub::RDO_TABLE_WORDS;
  if (lane < (1u << f.wb)) ub::rdo_fill_level(p, u, f, lane, table);
  __syncwarp();  // orders the lanes' table writes before lane 0's reads
  if (lane == 0) {
    ub::RdoCur cur;
    ub::rdo_prepare_fast(p, blk, f, table, prm, cur);
    pre[(size_t)c * RDO_CHUNK + k] = ub::rdo_pre_of(cur);
  }
}

// hazard[bi] = 1 when a block of a mode that reads a mode-0 block's endpoint bits follows within the window in the
// same chain (then a mode-0 endpoint re-fit at bi must happen inside the walk; otherwise it is postponed to k_rdo_rehint)
__global__ void __launch_bounds__(256) k_rdo_hazard(const uint4* __restrict__ blocks, const RdoChain* __restrict__ chains, uint32_t n_chains,
                                                    uint32_t window, uint8_t* __restrict__ hazard) {
  const uint32_t c = blockIdx.y;
  if (c >= n_chains) return;
  const RdoChain ch = chains[c];
  for (uint32_t bi = ch.first + blockIdx.x * blockDim.x + threadIdx.x; bi < ch.last; bi += gridDim.x * blockDim.x) {
    uint8_t hz = 0;
    if (ub::UBT(uastc_huff_modes)[blocks[bi].x & 127u] == 0) {
      const uint32_t end = (uint32_t)min((uint64_t)ch.last, (uint64_t)bi + 1 + window);
      for (uint32_t j = bi + 1; j < end; j++)
        if (ub::rdo_mode_reads_mode0_endpoints(ub::UBT(uastc_huff_modes)[blocks[j].x & 127u])) { hz = 1; break; }
    }
    hazard[bi] = hz;
  }
}

__global__ void __launch_bounds__(RDO_THREADS) k_rdo_chain(uint4* __restrict__ blocks, const uint4* __restrict__ px, const RdoChain* __restrict__ chains,
                                                           uint32_t step0, const ub::RdoBlockInfo* __restrict__ info,
                                                           const ub::RdoPre* __restrict__ pre, const uint32_t* __restrict__ tables,
                                                           ub::RdoHistEntry* __restrict__ hist_all, const uint8_t* __restrict__ hazard,
                                                           uint8_t* __restrict__ modified, ub::RdoParams p) {
  // per-step inputs are double-buffered: while step bi is scored, the loads of step bi+1 are in flight
  __shared__ __align__(16) uint32_t s_table[2][ub::RDO_TABLE_WORDS];
  __shared__ __align__(16) uint32_t s_px[2][16];
  __shared__ __align__(16) uint4 s_blk[2];
  __shared__ __align__(16) ub::RdoPre s_pre[2];
  __shared__ ub::RdoBlockInfo s_info[2];
  __shared__ __align__(16) uint4 s_ring[256];
  __shared__ float s_keep_t;
  __shared__ unsigned long long s_best;

  const RdoChain ch = chains[blockIdx.x];
  const uint32_t tid = threadIdx.x;
  ub::RdoHistEntry* hist = hist_all + ch.hist_off;
  const uint32_t begin = ch.first + step0;
  if (begin >= ch.last) return;
  const uint32_t end = min(ch.last, begin + RDO_CHUNK);
  // refill the ring with the (final) blocks just before this chunk
  if (tid < 256) { const uint32_t back = tid + 1; if (begin >= ch.first + back) s_ring[(begin - back) & 255u] = blocks[begin - back]; }

  // what each thread fetches for a step: 128 threads the table, 4 the texels, one each the block, info and pre-computed stats
  uint4 nx = make_uint4(0, 0, 0, 0);
  ub::RdoBlockInfo nx_info;
  ub::RdoPre nx_pre;
  auto fetch = [&](uint32_t bi) {
    const size_t slot = (size_t)blockIdx.x * RDO_CHUNK + (bi - begin);
    if (tid < 128) nx = __ldg((const uint4*)(tables + slot * ub::RDO_TABLE_WORDS) + tid);
    else if (tid < 132) nx = __ldg(px + (size_t)bi * 4 + (tid - 128));
    else if (tid == 132) nx = blocks[bi];
    else if (tid == 133) nx_info = info[slot];
    else if (tid == 134) nx_pre = pre[slot];
  };
  auto stash = [&](uint32_t b) {
    if (tid < 128) ((uint4*)s_table[b])[tid] = nx;
    else if (tid < 132) ((uint4*)s_px[b])[tid - 128] = nx;
    else if (tid == 132) s_blk[b] = nx;
    else if (tid == 133) s_info[b] = nx_info;
    else if (tid == 134) s_pre[b] = nx_pre;
  };
  fetch(begin);
  stash(0);
  if (tid == 0) s_best = ~0ull;
  __syncthreads();

  uint32_t b = 0;
  for (uint32_t bi = begin; bi < end; bi++, b ^= 1) {
    if (bi + 1 < end) fetch(bi + 1);  // block bi+1 is not rewritten before its own step, so reading it now is safe
    const ub::RdoBlockInfo f = s_info[b];
    const ub::Bits128 blk = to_bits(s_blk[b]);
    const bool active = f.valid && f.mode != ub::UASTC_MODE_SOLID;  // uniform; solid blocks are left alone but stay splice sources
    const uint32_t first = ub::UBT(rdo_sel_bits)[f.mode * 2], total = ub::UBT(rdo_sel_bits)[f.mode * 2 + 1];
    ub::RdoCur c;
    {
      const ub::RdoPre q = s_pre[b];
      c.mode = f.mode; c.first = first; c.total = total; c.sel_lo = q.sel_lo;
      c.ms_err = q.ms_err; c.rms_err = q.rms_err; c.smooth = q.smooth; c.solid = false; c.skip = q.skip != 0;
    }
    if (active && !c.skip) {
      const uint32_t lo_index = (bi - ch.first > p.window) ? bi - p.window : ch.first;
      const uint32_t n_win = bi - lo_index;
      // thread k takes candidates k, k+256, ... (k = 0 is the nearest earlier block); the spare warp prices "keep as is"
      if (tid == 256) s_keep_t = ub::rdo_keep_cost(c, ub::rdo_hist_find(hist, ch.hist_mask, first, c.sel_lo), bi, p);
      if (tid < 256) {
        unsigned long long my_key = ~0ull;
        for (uint32_t k = tid; k < n_win; k += 256) {
          const uint32_t pi = bi - 1 - k;
          const ub::Bits128 prev = (k < 256) ? to_bits(s_ring[pi & 255u]) : to_bits(blocks[pi]);  // window beyond the ring: global
          uint64_t slo, shi;
          ub::rdo_get_sel(prev, first, total, slo, shi);
          const int64_t h = ub::rdo_hist_find(hist, ch.hist_mask, first, slo);
          const uint32_t match = h >= 0 ? (uint32_t)h : pi;
          if (match > pi) continue;  // a later block already carries these bits: it was (or will be) tried instead
          float t;
          if (!ub::rdo_trial_fast(s_table[b], f, slo, shi, c, bi, match, p, t)) continue;
          const unsigned long long key = ((unsigned long long)__float_as_uint(t) << 32) | k;  // t >= 0: bit order == value order
          my_key = key < my_key ? key : my_key;
        }
        if (my_key != ~0ull) atomicMin(&s_best, my_key);
      }
      __syncthreads();
    }
    if (tid == 0) {
      ub::Bits128 best = blk;
      if (active) {
        uint64_t key_lo = c.sel_lo;
        if (!c.skip) {
          const unsigned long long w = s_best;
          s_best = ~0ull;
          if (w != ~0ull && __uint_as_float((uint32_t)(w >> 32)) < s_keep_t) {
            const uint32_t pi = bi - 1 - (uint32_t)(w & 0xFFFFFFFFu);
            const ub::Bits128 prev = (bi - pi <= 256) ? to_bits(s_ring[pi & 255u]) : to_bits(blocks[pi]);
            best = ub::rdo_splice(blk, prev, first, total);
            uint8_t mark = 1;
            if (p.refine && f.mode == 0) {
              if (hazard[bi]) best = ub::rdo_refine_mode0(s_px[b], best);  // a later block will read these endpoint bits
              else mark = 2;                                               // re-fit postponed to k_rdo_rehint
            }
            blocks[bi] = from_bits(best);
            modified[bi] = mark;
            uint64_t hi;
            ub::rdo_get_sel(best, first, total, key_lo, hi);
          }
        }
        ub::rdo_hist_set(hist, ch.hist_mask, first, key_lo, bi);
      }
      s_ring[bi & 255u] = from_bits(best);
    }
    if (bi + 1 < end) stash(b ^ 1);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(128) k_rdo_rehint(uint4* __restrict__ blocks, const uint4* __restrict__ px, const uint8_t* __restrict__ modified,
                                                    uint64_t n, uint32_t flags) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    if (!modified[i]) continue;
    uint32_t p[16];
#pragma unroll
    for (uint32_t y = 0; y < 4; y++) {
      const uint4 v = __ldg(px + i * 4 + y);
      p[4 * y] = v.x; p[4 * y + 1] = v.y; p[4 * y + 2] = v.z; p[4 * y + 3] = v.w;
    }
    ub::Bits128 b = to_bits(blocks[i]);
    if (modified[i] == 2) b = ub::rdo_refine_mode0(p, b);
    blocks[i] = from_bits(ub::rdo_rehint(p, b, flags));
  }
}

const int MAXC_FAST = 21, MAXC_FULL = 170;

// Encode launch shape: 1024-thread CTAs (one per SM, every warp of the SM in the same phase) when there is enough
// work to give every SM two of them; smaller CTAs for small inputs so that all SMs still get blocks.
unsigned encode_threads(uint64_t n) {
  unsigned t = ENC_THREADS;
  while (t > 128 && (n + t - 1) / t < 2ull * (uint64_t)g_ctx.sm_count) t >>= 1;
  return t;
}
unsigned encode_grid(uint64_t n) {
  const unsigned t = encode_threads(n);
  const uint64_t want = (n + t - 1) / t;
  return (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(want, (uint64_t)g_ctx.sm_count * 64));
}

__global__ void k_set_u64(unsigned long long* p, unsigned long long v) { *p = v; }

int launch_transcode(uint32_t target, const void* d_src, void* d_dst, uint64_t n, cudaStream_t st, uint64_t* d_first_bad) {
  if (!n) return B200_OK;
  // persistent tiles: up to 6 CTAs (34 KB of shared memory each) per SM, each walking tiles with a grid stride
  const uint64_t tiles = (n + XT_TILE - 1) / XT_TILE;
  const unsigned grid = (unsigned)std::min<uint64_t>(tiles, (uint64_t)g_ctx.sm_count * XT_CTAS_PER_SM);
  if (target == B200_TTF_BC7_RGBA)
    k_transcode_tiled<0><<<grid, XT_THREADS, xt_smem_bytes(0), st>>>((const uint4*)d_src, (uint4*)d_dst, n, (unsigned long long*)d_first_bad, XtImage{});
  else
    k_transcode_tiled<1><<<grid, XT_THREADS, xt_smem_bytes(1), st>>>((const uint4*)d_src, (uint4*)d_dst, n, (unsigned long long*)d_first_bad, XtImage{});
  g_launches++;
  CU(cudaGetLastError());
  return B200_OK;
}

// The hint-driven block targets (ETC1_RGB 0, ETC2_RGBA 1, BC1_RGB 2, BC3_RGBA 3, BC4_R 4, BC5_RG 5): 8 or 16 bytes per block.
bool is_hinted_target(uint32_t t) { return t <= B200_TTF_BC5_RG || t == B200_TTF_ETC2_EAC_R11 || t == B200_TTF_ETC2_EAC_RG11; }
uint32_t out_block_bytes(uint32_t t) { return (t == B200_TTF_ETC1_RGB || t == B200_TTF_BC1_RGB || t == B200_TTF_BC4_R || t == B200_TTF_ETC2_EAC_R11) ? 8u : 16u; }
bool is_block_target(uint32_t t) { return t == B200_TTF_BC7_RGBA || t == B200_TTF_ASTC_4x4_RGBA || is_hinted_target(t); }

int launch_hinted(uint32_t target, uint32_t decode_flags, uint32_t has_alpha, const void* d_src, void* d_dst, uint64_t n, cudaStream_t st,
                  uint64_t* d_first_bad) {
  if (!n) return B200_OK;
  const bool from_alpha = has_alpha && (decode_flags & B200_DECODE_FLAG_ALPHA_TO_OPAQUE);
  if (target == B200_TTF_ETC1_RGB && from_alpha) return fail(B200_E_INVALID_VALUE, "ETC1 from the alpha channel is not accelerated");
  if (((uintptr_t)d_dst & 15u) || ((uintptr_t)d_src & 15u)) return fail(B200_E_INVALID_VALUE, "device pointers must be 16-byte aligned");
  XtImage im{};
  im.kind = target;
  im.hq = (decode_flags & B200_DECODE_FLAG_HIGH_QUALITY) ? 1u : 0u;
  im.chan0 = ((target == B200_TTF_BC4_R || target == B200_TTF_ETC2_EAC_R11) && from_alpha) ? 3u : 0u;   // transcode_image, XC:10417-10438, :10515
  im.chan1 = 3u;
  const uint64_t tiles = (n + xt_tile(3) - 1) / xt_tile(3);
  const unsigned grid = (unsigned)std::min<uint64_t>(tiles, (uint64_t)g_ctx.sm_count * XT_CTAS_PER_SM);
  switch (target == B200_TTF_ETC2_EAC_R11 ? 6u : target == B200_TTF_ETC2_EAC_RG11 ? 7u : target) {
#define XH_CASE(KIND) case KIND: k_transcode_tiled<3, KIND><<<grid, XT_THREADS, xt_smem_bytes(3), st>>>((const uint4*)d_src, (uint4*)d_dst, n, (unsigned long long*)d_first_bad, im); break;
    XH_CASE(0) XH_CASE(1) XH_CASE(2) XH_CASE(3) XH_CASE(4) XH_CASE(5) XH_CASE(6) XH_CASE(7)
#undef XH_CASE
  }
  g_launches++;
  CU(cudaGetLastError());
  return B200_OK;
}

bool is_texel_target(uint32_t t) { return t == B200_TTF_RGBA32 || t == B200_TTF_RGB565 || t == B200_TTF_BGR565 || t == B200_TTF_RGBA4444; }
uint32_t texel_bytes(uint32_t t) { return t == B200_TTF_RGBA32 ? 4u : 2u; }

// One image's blocks (raster order, ceil(w/4) x ceil(h/4)) -> w x h texels at d_dst, row pitch in texels.
int launch_decode(uint32_t target, const void* d_src, void* d_dst, uint32_t w, uint32_t h, uint32_t pitch_px, cudaStream_t st, uint64_t* d_first_bad) {
  if (!w || !h) return B200_OK;
  XtImage im;
  im.w = w; im.h = h; im.bx = (w + 3) / 4; im.pitch_px = pitch_px; im.kind = target;
  const uint32_t row_align = target == B200_TTF_RGBA32 ? 4u : 4u;   // 4 texels = 16 B (RGBA32) or 8 B (16-bit targets)
  im.vec_ok = (pitch_px % row_align == 0 && ((uintptr_t)d_dst % (target == B200_TTF_RGBA32 ? 16 : 8)) == 0) ? 1u : 0u;
  const uint64_t n = (uint64_t)im.bx * ((h + 3) / 4);
  const uint64_t tiles = (n + xt_tile(2) - 1) / xt_tile(2);
  const unsigned grid = (unsigned)std::min<uint64_t>(tiles, (uint64_t)g_ctx.sm_count * XT_CTAS_PER_SM);
  k_transcode_tiled<2><<<grid, XT_THREADS, xt_smem_bytes(2), st>>>((const uint4*)d_src, (uint4*)d_dst, n, (unsigned long long*)d_first_bad, im);
  g_launches++;
  CU(cudaGetLastError());
  return B200_OK;
}

// The RDO post-pass over n device-resident blocks that form `n_images` consecutive images (image i holds
// img_blocks[i] blocks; job ranges never cross an image, ENC/basisu_comp.cpp:1574 calls uastc_rdo per slice).
int rdo_run(uint4* d_blocks, const uint4* d_px, const std::vector<uint64_t>& img_blocks, const b200_uastc_params& prm, cudaStream_t st) {
  ub::RdoParams p;
  p.lambda = prm.rdo_lambda > 0.0f ? prm.rdo_lambda : 1.0f;
  p.max_rms_ratio = 10.0f;   // UASTC_RDO_DEFAULT_MAX_ALLOWED_RMS_INCREASE_RATIO (libktx does not expose it)
  p.skip_rms = 8.0f;         // UASTC_RDO_DEFAULT_SKIP_BLOCK_RMS_THRESH
  p.smooth_scale = prm.rdo_max_smooth_block_error_scale > 0.0f ? prm.rdo_max_smooth_block_error_scale : 10.0f;
  p.smooth_std_dev = prm.rdo_max_smooth_block_std_dev > 0.0f ? prm.rdo_max_smooth_block_std_dev : 18.0f;
  const uint32_t dict = prm.rdo_dict_size ? prm.rdo_dict_size : 4096u;
  p.window = std::max<uint32_t>(1u, dict / 16u);
  p.literal_cost = 100; p.flags = prm.pack_flags; p.refine = 1;

  std::vector<RdoChain> chains;
  uint64_t base = 0, hist_total = 0, n_total = 0;
  uint32_t longest = 0;
  for (uint64_t nb : img_blocks) {
    uint32_t per;
    const uint32_t jobs = ub::rdo_job_ranges((uint32_t)nb, prm.rdo_total_jobs, &per);
    for (uint32_t j = 0; j < jobs; j++) {
      RdoChain c;
      c.first = (uint32_t)(base + (uint64_t)j * per);
      c.last = (uint32_t)(base + std::min<uint64_t>(nb, (uint64_t)(j + 1) * per));
      uint32_t cap = 16;
      while (cap < 2 * (c.last - c.first) + 16) cap <<= 1;
      c.hist_off = hist_total; c.hist_mask = cap - 1; c.pad = 0;
      hist_total += cap;
      longest = std::max(longest, c.last - c.first);
      chains.push_back(c);
    }
    base += nb;
  }
  n_total = base;
  if (n_total >= 0xFFFFFFFFull) return fail(B200_E_INVALID_VALUE, "too many blocks for one RDO call");
  if (chains.empty()) return B200_OK;
  const uint32_t nc = (uint32_t)chains.size();

  RdoChain* d_chains = nullptr; ub::RdoHistEntry* d_hist = nullptr; ub::RdoBlockInfo* d_info = nullptr;
  uint32_t* d_tables = nullptr; uint8_t* d_mod = nullptr; unsigned int* d_err = nullptr; ub::RdoPre* d_pre = nullptr; uint8_t* d_hz = nullptr;
  auto cleanup = [&]() { cudaFree(d_chains); cudaFree(d_hist); cudaFree(d_info); cudaFree(d_tables); cudaFree(d_mod); cudaFree(d_err); cudaFree(d_pre); cudaFree(d_hz); };
  cudaError_t e = cudaSuccess;
  if (e == cudaSuccess) e = cudaMalloc(&d_chains, nc * sizeof(RdoChain));
  if (e == cudaSuccess) e = cudaMalloc(&d_hist, hist_total * sizeof(ub::RdoHistEntry));
  if (e == cudaSuccess) e = cudaMalloc(&d_info, (size_t)nc * RDO_CHUNK * sizeof(ub::RdoBlockInfo));
  if (e == cudaSuccess) e = cudaMalloc(&d_tables, (size_t)nc * RDO_CHUNK * ub::RDO_TABLE_WORDS * 4);
  if (e == cudaSuccess) e = cudaMalloc(&d_pre, (size_t)nc * RDO_CHUNK * sizeof(ub::RdoPre));
  if (e == cudaSuccess) e = cudaMalloc(&d_hz, n_total);
  if (e == cudaSuccess) e = cudaMalloc(&d_mod, n_total);
  if (e == cudaSuccess) e = cudaMalloc(&d_err, 4);
  if (e != cudaSuccess) { cleanup(); return fail(B200_E_OUT_OF_MEMORY, "cudaMalloc RDO state", e); }
  cudaMemcpyAsync(d_chains, chains.data(), nc * sizeof(RdoChain), cudaMemcpyHostToDevice, st);
  cudaMemsetAsync(d_hist, 0, hist_total * sizeof(ub::RdoHistEntry), st);
  cudaMemsetAsync(d_mod, 0, n_total, st);
  cudaMemsetAsync(d_err, 0, 4, st);
  k_rdo_hazard<<<dim3((unsigned)std::min<uint32_t>((longest + 255) / 256, 1024u), nc), 256, 0, st>>>(d_blocks, d_chains, nc, p.window, d_hz);
  g_launches++;
  for (uint32_t step0 = 0; step0 < longest; step0 += RDO_CHUNK) {
    const uint64_t warps = (uint64_t)nc * RDO_CHUNK;
    k_rdo_prep<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, st>>>(d_blocks, d_px, d_chains, nc, step0, d_info, d_pre, d_tables, d_err, p);
    k_rdo_chain<<<nc, RDO_THREADS, 0, st>>>(d_blocks, d_px, d_chains, step0, d_info, d_pre, d_tables, d_hist, d_hz, d_mod, p);
    g_launches += 2;
  }
  k_rdo_rehint<<<(unsigned)std::min<uint64_t>((n_total + 127) / 128, (uint64_t)g_ctx.sm_count * 64), 128, 0, st>>>(d_blocks, d_px, d_mod, n_total, prm.pack_flags);
  g_launches++;
  unsigned int h_err = 0;
  e = cudaMemcpyAsync(&h_err, d_err, 4, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  cleanup();
  if (e != cudaSuccess) return fail(B200_E_CUDA, "RDO pipeline", e);
  if (h_err) return fail(B200_E_BAD_BLOCK, "RDO: an input block does not unpack");
  return B200_OK;
}

// Encode + RDO for a host call: every image's texel blocks and UASTC blocks stay resident (64 + 16 B per block) so the
// post-pass can walk them; job ranges are per image, all ranges of all images walk concurrently.
unsigned expand_grid(uint64_t texels) { return (unsigned)std::min<uint64_t>((texels + 255) / 256, (uint64_t)g_ctx.sm_count * 32); }

// Source layout of a host encode call: components per texel and the packed component mapping (see k_expand_swizzle)
struct SrcFmt { uint32_t comps = 4, sel = 0x04030201u; bool identity() const { return comps == 4 && sel == 0x04030201u; } };

int encode_with_rdo(const b200_image* images, uint32_t n_images, const SrcFmt& fmt, const b200_uastc_params* params, uint8_t* const* out_blocks,
                    uint32_t* out_any_alpha, cudaStream_t* ss) {
  std::vector<uint64_t> nb(n_images);
  uint64_t total = 0;
  size_t max_in = 0, max_exp = 0;
  for (uint32_t i = 0; i < n_images; i++) {
    const size_t pitch = images[i].row_pitch_bytes ? images[i].row_pitch_bytes : (size_t)images[i].width * fmt.comps;
    if (pitch < (size_t)images[i].width * fmt.comps) return fail(B200_E_INVALID_VALUE, "row pitch smaller than a row");
    nb[i] = (uint64_t)((images[i].width + 3) / 4) * ((images[i].height + 3) / 4);
    total += nb[i];
    max_in = std::max(max_in, pitch * images[i].height);
    max_exp = std::max(max_exp, (size_t)images[i].width * 4 * images[i].height);
  }
  uint8_t* d_img = nullptr; uint4* d_px = nullptr; uint4* d_blk = nullptr; unsigned int* d_alpha = nullptr; uint8_t* d_exp = nullptr;
  auto cleanup = [&]() { cudaFree(d_img); cudaFree(d_px); cudaFree(d_blk); cudaFree(d_alpha); cudaFree(d_exp); };
  cudaError_t e = cudaMalloc(&d_img, max_in);
  if (e == cudaSuccess && !fmt.identity()) e = cudaMalloc(&d_exp, max_exp);
  if (e == cudaSuccess) e = cudaMalloc(&d_px, total * 64);
  if (e == cudaSuccess) e = cudaMalloc(&d_blk, total * 16);
  if (e == cudaSuccess) e = cudaMalloc(&d_alpha, 4);
  if (e != cudaSuccess) { cleanup(); return fail(B200_E_OUT_OF_MEMORY, "cudaMalloc RDO buffers", e); }
  cudaStream_t st = ss[1];
  cudaMemsetAsync(d_alpha, 0, 4, st);
  const bool full = (params->pack_flags & 7u) >= 4u;
  uint64_t off = 0;
  for (uint32_t i = 0; i < n_images && e == cudaSuccess; i++) {
    const b200_image& im = images[i];
    uint32_t pitch = im.row_pitch_bytes ? im.row_pitch_bytes : im.width * fmt.comps;
    const uint32_t bx = (im.width + 3) / 4, by = (im.height + 3) / 4;
    e = cudaMemcpyAsync(d_img, im.rgba, (size_t)pitch * im.height, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) break;
    const uint8_t* d_rgba = d_img;
    if (!fmt.identity()) {
      k_expand_swizzle<<<expand_grid((uint64_t)im.width * im.height), 256, 0, st>>>(d_img, im.width, im.height, pitch, fmt.comps, fmt.sel, (uint32_t*)d_exp);
      g_launches++;
      d_rgba = d_exp; pitch = im.width * 4;
    }
    k_gather_blocks<<<(unsigned)std::min<uint64_t>((nb[i] + 255) / 256, (uint64_t)g_ctx.sm_count * 16), 256, 0, st>>>(d_rgba, im.width, im.height, pitch, d_px + off * 4, bx, by, d_alpha);
    if (full) k_encode_blocks<MAXC_FULL><<<encode_grid(nb[i]), encode_threads(nb[i]), 0, st>>>(d_px + off * 4, d_blk + off, nb[i], params->pack_flags);
    else k_encode_blocks<MAXC_FAST><<<encode_grid(nb[i]), encode_threads(nb[i]), 0, st>>>(d_px + off * 4, d_blk + off, nb[i], params->pack_flags);
    g_launches += 2;
    off += nb[i];
  }
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) { cleanup(); return fail(B200_E_CUDA, "encode before RDO", e); }
  int rc = rdo_run(d_blk, d_px, nb, *params, st);
  if (rc) { cleanup(); return rc; }
  off = 0;
  for (uint32_t i = 0; i < n_images && e == cudaSuccess; i++) {
    e = cudaMemcpyAsync(out_blocks[i], d_blk + off, nb[i] * 16, cudaMemcpyDeviceToHost, st);
    off += nb[i];
  }
  unsigned int h_alpha = 0;
  if (e == cudaSuccess) e = cudaMemcpyAsync(&h_alpha, d_alpha, 4, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cleanup();
  if (e != cudaSuccess) return fail(B200_E_CUDA, "RDO download", e);
  if (out_any_alpha) *out_any_alpha = h_alpha ? 1u : 0u;
  return B200_OK;
}

// per-thread stream set, created lazily
struct ThreadStreams {
  cudaStream_t s[4] = {nullptr, nullptr, nullptr, nullptr};   // copy in, kernels, copy out, second kernel stream
  int device = -1;
  ~ThreadStreams() { /* streams die with the context */ }
};
thread_local ThreadStreams t_streams;

// Per host thread: the staging buffers of the host-buffer entry points live across calls (a cudaMalloc / cudaFree pair per
// call costs milliseconds and synchronises the device).  Slots 0-2: slab inputs, 3-5: slab outputs, 6: flags.
struct Staging {
  void* p[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // 7-9: expanded RGBA8 slabs
  size_t n[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  void* get(int i, size_t bytes) {
    if (n[i] < bytes) {
      if (p[i]) cudaFree(p[i]);
      p[i] = nullptr; n[i] = 0;
      if (cudaMalloc(&p[i], bytes) != cudaSuccess) return nullptr;
      n[i] = bytes;
    }
    return p[i];
  }
  void release() { for (int i = 0; i < 10; i++) { if (p[i]) cudaFree(p[i]); p[i] = nullptr; n[i] = 0; } }
  ~Staging() { /* at thread exit the context may be gone: the driver reclaims */ }
};
thread_local Staging t_stage;

int get_streams(cudaStream_t** out) {
  if (t_streams.device != g_ctx.device) {
    CU(cudaSetDevice(g_ctx.device));
    for (int i = 0; i < 4; i++) CU(cudaStreamCreateWithFlags(&t_streams.s[i], cudaStreamNonBlocking));
    t_streams.device = g_ctx.device;
  }
  *out = t_streams.s;
  return B200_OK;
}

}  // namespace

extern "C" {

int b200_uastc_init(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_ctx.ready) return B200_OK;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) return fail(B200_E_NO_DEVICE, "no CUDA device (this library has no CPU fallback)", e);
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  cudaDeviceProp p;
  CU(cudaGetDeviceProperties(&p, dev));
  if (p.major < 10) return fail(B200_E_NO_DEVICE, "device is not sm_100 (B200) class; kernels are built for sm_100a only");
  CU(cudaSetDevice(dev));
  g_ctx.device = dev;
  g_ctx.sm_count = p.multiProcessorCount;
  CU(cudaFuncSetAttribute(k_transcode_tiled<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(0)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(1)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(2)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<3, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(3)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<3, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(3)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(3)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<3, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(3)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<3, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(3)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<3, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(3)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<3, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(3)));
  CU(cudaFuncSetAttribute(k_transcode_tiled<3, 7>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xt_smem_bytes(3)));
  g_ctx.ready = true;
  return B200_OK;
}

void b200_uastc_shutdown(void) {
  t_stage.release();
  std::lock_guard<std::mutex> lk(g_mu);
  g_ctx.ready = false;
}

const char* b200_uastc_last_error(void) { return g_err.c_str(); }
uint64_t b200_uastc_launch_count(void) { return g_launches.load(); }

int b200_uastc_transcode_dev(uint32_t target, const void* d_src, void* d_dst, uint64_t n_blocks, void* stream, uint64_t* d_first_bad) {
  return b200_uastc_transcode_dev_ex(target, 0, 1, d_src, d_dst, n_blocks, stream, d_first_bad);
}

int b200_uastc_transcode_dev_ex(uint32_t target, uint32_t decode_flags, uint32_t has_alpha, const void* d_src, void* d_dst, uint64_t n_blocks,
                                void* stream, uint64_t* d_first_bad) {
  if (!is_block_target(target)) return fail(B200_E_INVALID_VALUE, "unsupported transcode target");
  if (n_blocks && (!d_src || !d_dst)) return fail(B200_E_INVALID_VALUE, "null device pointer");
  int rc = b200_uastc_init();
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (!st) { cudaStream_t* ss; rc = get_streams(&ss); if (rc) return rc; st = ss[0]; }
  if (d_first_bad) { k_set_u64<<<1, 1, 0, st>>>((unsigned long long*)d_first_bad, ~0ull); g_launches++; }
  rc = is_hinted_target(target) ? launch_hinted(target, decode_flags, has_alpha, d_src, d_dst, n_blocks, st, d_first_bad)
                                : launch_transcode(target, d_src, d_dst, n_blocks, st, d_first_bad);
  if (rc) return rc;
  if (!stream) CU(cudaStreamSynchronize(st));
  return B200_OK;
}

int b200_uastc_decode_image_dev(uint32_t target, const void* d_src, void* d_dst, uint32_t width, uint32_t height,
                                uint32_t dst_row_pitch_texels, void* stream, uint64_t* d_first_bad) {
  if (!is_texel_target(target)) return fail(B200_E_INVALID_VALUE, "unsupported decode target");
  if (width && height && (!d_src || !d_dst)) return fail(B200_E_INVALID_VALUE, "null device pointer");
  if (dst_row_pitch_texels && dst_row_pitch_texels < width) return fail(B200_E_INVALID_VALUE, "row pitch smaller than the image");
  if (((uintptr_t)d_dst % texel_bytes(target)) != 0) return fail(B200_E_INVALID_VALUE, "destination not aligned to the texel size");
  int rc = b200_uastc_init();
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (!st) { cudaStream_t* ss; rc = get_streams(&ss); if (rc) return rc; st = ss[0]; }
  if (d_first_bad) { k_set_u64<<<1, 1, 0, st>>>((unsigned long long*)d_first_bad, ~0ull); g_launches++; }
  rc = launch_decode(target, d_src, d_dst, width, height, dst_row_pitch_texels ? dst_row_pitch_texels : width, st, d_first_bad);
  if (rc) return rc;
  if (!stream) CU(cudaStreamSynchronize(st));
  return B200_OK;
}

int b200_uastc_decode(uint32_t target, const b200_decode_image* images, uint32_t n_images, const uint8_t* src_base,
                      uint8_t* dst_base, int64_t* bad_image) {
  if (bad_image) *bad_image = -1;
  if (!is_texel_target(target)) return fail(B200_E_INVALID_VALUE, "unsupported decode target");
  if (n_images && (!images || !src_base || !dst_base)) return fail(B200_E_INVALID_VALUE, "null pointer");
  int rc = b200_uastc_init();
  if (rc) return rc;
  cudaStream_t* ss;
  rc = get_streams(&ss);
  if (rc) return rc;
  const uint32_t bpp = texel_bytes(target);

  // Slabs = whole block rows of one image: contiguous in the UASTC source and (rows of w texels) in the destination.
  struct Slab { uint32_t image, w, rows_px; uint64_t so, dofs, n_blocks; };
  std::vector<Slab> slabs;
  const uint64_t SLAB_BLOCKS = 1u << 19;   // 8 MiB in, up to 32 MiB out
  uint64_t max_in = 0, max_out = 0;
  for (uint32_t i = 0; i < n_images; i++) {
    const uint32_t w = images[i].width, h = images[i].height;
    if (!w || !h) continue;
    const uint32_t bx = (w + 3) / 4, by = (h + 3) / 4;
    const uint32_t rows_per = (uint32_t)std::max<uint64_t>(1, SLAB_BLOCKS / bx);
    for (uint32_t r0 = 0; r0 < by; r0 += rows_per) {
      const uint32_t nr = std::min(rows_per, by - r0);
      const uint32_t px_rows = std::min(h - r0 * 4, nr * 4);
      slabs.push_back({i, w, px_rows, images[i].src_offset + (uint64_t)r0 * bx * 16, images[i].dst_offset + (uint64_t)r0 * 4 * w * bpp, (uint64_t)nr * bx});
      max_in = std::max(max_in, (uint64_t)nr * bx * 16);
      max_out = std::max(max_out, (uint64_t)px_rows * w * bpp);
    }
  }
  if (slabs.empty()) return B200_OK;
  const int RING = 3;
  const int nring = (int)std::min<size_t>(RING, slabs.size());
  uint8_t* d_in[RING] = {0}; uint8_t* d_out[RING] = {0};
  cudaEvent_t ev_in[RING], ev_k[RING], ev_out[RING];
  for (int i = 0; i < nring; i++) {
    d_in[i] = (uint8_t*)t_stage.get(i, max_in); d_out[i] = (uint8_t*)t_stage.get(3 + i, max_out);
    if (!d_in[i] || !d_out[i]) return fail(B200_E_OUT_OF_MEMORY, "cudaMalloc staging");
  }
  unsigned long long* d_bad = (unsigned long long*)t_stage.get(6, (size_t)n_images * sizeof(unsigned long long));
  if (!d_bad) return fail(B200_E_OUT_OF_MEMORY, "cudaMalloc flags");
  for (int i = 0; i < nring; i++) {
    cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ev_k[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ev_out[i], cudaEventDisableTiming);
  }
  cudaStream_t s_in = ss[0], s_k = ss[1], s_out = ss[2];
  cudaMemsetAsync(d_bad, 0xFF, (size_t)n_images * sizeof(unsigned long long), s_k);
  cudaError_t err = cudaSuccess;
  for (size_t k = 0; k < slabs.size() && err == cudaSuccess; k++) {
    const Slab& sl = slabs[k];
    const int b = (int)(k % nring);
    if (k >= (size_t)nring) {
      cudaStreamWaitEvent(s_in, ev_k[b], 0);
      cudaStreamWaitEvent(s_k, ev_out[b], 0);
    }
    err = cudaMemcpyAsync(d_in[b], src_base + sl.so, sl.n_blocks * 16, cudaMemcpyHostToDevice, s_in);
    if (err != cudaSuccess) break;
    cudaEventRecord(ev_in[b], s_in);
    cudaStreamWaitEvent(s_k, ev_in[b], 0);
    launch_decode(target, d_in[b], d_out[b], sl.w, sl.rows_px, sl.w, s_k, (uint64_t*)(d_bad + sl.image));
    cudaEventRecord(ev_k[b], s_k);
    cudaStreamWaitEvent(s_out, ev_k[b], 0);
    err = cudaMemcpyAsync(dst_base + sl.dofs, d_out[b], (size_t)sl.rows_px * sl.w * bpp, cudaMemcpyDeviceToHost, s_out);
    cudaEventRecord(ev_out[b], s_out);
  }
  std::vector<unsigned long long> h_bad(n_images, ~0ull);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_in);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_k);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_out);
  if (err == cudaSuccess) err = cudaMemcpy(h_bad.data(), d_bad, (size_t)n_images * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  for (int i = 0; i < nring; i++) { cudaEventDestroy(ev_in[i]); cudaEventDestroy(ev_k[i]); cudaEventDestroy(ev_out[i]); }
  if (err != cudaSuccess) return fail(B200_E_CUDA, "decode pipeline", err);
  for (uint32_t i = 0; i < n_images; i++)
    if (h_bad[i] != ~0ull) {
      if (bad_image) *bad_image = i;
      return fail(B200_E_BAD_BLOCK, "invalid UASTC block");
    }
  return B200_OK;
}

int b200_uastc_transcode(uint32_t target, const b200_xcode_span* spans, uint32_t n_spans, const uint8_t* src_base,
                         uint8_t* dst_base, int64_t* bad_span) {
  return b200_uastc_transcode_ex(target, 0, 1, spans, n_spans, src_base, dst_base, bad_span);
}

int b200_uastc_transcode_ex(uint32_t target, uint32_t decode_flags, uint32_t has_alpha, const b200_xcode_span* spans, uint32_t n_spans,
                            const uint8_t* src_base, uint8_t* dst_base, int64_t* bad_span) {
  if (bad_span) *bad_span = -1;
  if (!is_block_target(target)) return fail(B200_E_INVALID_VALUE, "unsupported transcode target");
  const bool hinted = is_hinted_target(target);
  if (hinted && target == B200_TTF_ETC1_RGB && has_alpha && (decode_flags & B200_DECODE_FLAG_ALPHA_TO_OPAQUE))
    return fail(B200_E_INVALID_VALUE, "ETC1 from the alpha channel is not accelerated");
  const uint64_t ob = out_block_bytes(target);
  if (n_spans && (!spans || !src_base || !dst_base)) return fail(B200_E_INVALID_VALUE, "null pointer");
  int rc = b200_uastc_init();
  if (rc) return rc;
  cudaStream_t* ss;
  rc = get_streams(&ss);
  if (rc) return rc;

  // Slab list: spans are cut into <= SLAB blocks; adjacent spans that are contiguous in both src and dst merge.
  struct Run { uint64_t so, dofs, n; uint32_t first_span; };
  std::vector<Run> runs;
  for (uint32_t i = 0; i < n_spans; i++) {
    if (!spans[i].n_blocks) continue;
    // (the hint-driven targets keep one run -- one flag -- per span: a valid BC4 / BC5 block can be all zero, so the
    // zero-block scan below that names the span inside a merged run does not apply to them)
    if (!runs.empty() && !hinted) {
      Run& r = runs.back();
      if (r.so + r.n * 16 == spans[i].src_offset && r.dofs + r.n * ob == spans[i].dst_offset) { r.n += spans[i].n_blocks; continue; }
    }
    runs.push_back({spans[i].src_offset, spans[i].dst_offset, spans[i].n_blocks, i});
  }
  if (runs.empty()) return B200_OK;

  const uint64_t SLAB = 1u << 20;  // 16 MiB in, 16 MiB out per slab
  const int RING = 3;
  uint8_t* d_in[RING] = {0}; uint8_t* d_out[RING] = {0};
  unsigned long long* d_bad = nullptr;
  cudaEvent_t ev_in[RING], ev_k[RING], ev_out[RING];
  uint64_t maxn = 0;
  for (auto& r : runs) maxn = std::max(maxn, std::min(r.n, SLAB));
  auto cleanup = [&]() {};   // staging buffers stay cached (t_stage)
  int nring = 0;
  uint64_t total_slabs = 0;
  for (auto& r : runs) total_slabs += (r.n + SLAB - 1) / SLAB;
  nring = (int)std::min<uint64_t>(RING, total_slabs);
  for (int i = 0; i < nring; i++) {
    d_in[i] = (uint8_t*)t_stage.get(i, maxn * 16); d_out[i] = (uint8_t*)t_stage.get(3 + i, maxn * ob);
    if (!d_in[i] || !d_out[i]) return fail(B200_E_OUT_OF_MEMORY, "cudaMalloc staging");
    cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ev_k[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ev_out[i], cudaEventDisableTiming);
  }
  // one "first bad block" word per run so the failing span can be named
  d_bad = (unsigned long long*)t_stage.get(6, runs.size() * sizeof(unsigned long long));
  if (!d_bad) return fail(B200_E_OUT_OF_MEMORY, "cudaMalloc flags");
  cudaMemsetAsync(d_bad, 0xFF, runs.size() * sizeof(unsigned long long), ss[1]);

  cudaStream_t s_in = ss[0], s_k = ss[1], s_out = ss[2];
  uint64_t slab_idx = 0;
  cudaError_t err = cudaSuccess;
  for (size_t ri = 0; ri < runs.size() && err == cudaSuccess; ri++) {
    const Run& r = runs[ri];
    for (uint64_t o = 0; o < r.n && err == cudaSuccess; o += SLAB, slab_idx++) {
      const uint64_t n = std::min(SLAB, r.n - o);
      const int b = (int)(slab_idx % nring);
      if (slab_idx >= (uint64_t)nring) {
        cudaStreamWaitEvent(s_in, ev_k[b], 0);    // input buffer free once its kernel ran
        cudaStreamWaitEvent(s_k, ev_out[b], 0);   // output buffer free once copied back
      }
      err = cudaMemcpyAsync(d_in[b], src_base + r.so + o * 16, n * 16, cudaMemcpyHostToDevice, s_in);
      if (err != cudaSuccess) break;
      cudaEventRecord(ev_in[b], s_in);
      cudaStreamWaitEvent(s_k, ev_in[b], 0);
      // block indices inside the kernel are slab-relative; the flag only has to say "this run holds a bad block"
      if (hinted) launch_hinted(target, decode_flags, has_alpha, d_in[b], d_out[b], n, s_k, (uint64_t*)(d_bad + ri));
      else launch_transcode(target, d_in[b], d_out[b], n, s_k, (uint64_t*)(d_bad + ri));
      cudaEventRecord(ev_k[b], s_k);
      cudaStreamWaitEvent(s_out, ev_k[b], 0);
      err = cudaMemcpyAsync(dst_base + r.dofs + o * ob, d_out[b], n * ob, cudaMemcpyDeviceToHost, s_out);
      cudaEventRecord(ev_out[b], s_out);
    }
  }
  std::vector<unsigned long long> h_bad(runs.size(), ~0ull);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_in);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_k);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_out);
  if (err == cudaSuccess) err = cudaMemcpy(h_bad.data(), d_bad, runs.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  for (int i = 0; i < nring; i++) { cudaEventDestroy(ev_in[i]); cudaEventDestroy(ev_k[i]); cudaEventDestroy(ev_out[i]); }
  cleanup();
  if (err != cudaSuccess) return fail(B200_E_CUDA, "transcode pipeline", err);

  for (size_t ri = 0; ri < runs.size(); ri++) {
    if (h_bad[ri] != ~0ull) {
      // A run may hold several merged spans and several slabs; the flag only says "this run has a bad block".
      // Name the first span of the run that contains one by rescanning that run's spans on the device output:
      // cheap path -- report the run's first span unless the run was merged, in which case walk its spans.
      if (bad_span) {
        int64_t which = runs[ri].first_span;
        // refine: spans of this run are contiguous; a bad block's output is all-zero, which no valid BC7/ASTC
        // block produced here can be (BC7 mode bit / ASTC block mode are never zero).
        uint64_t off = runs[ri].dofs;
        for (uint32_t si = runs[ri].first_span; si < n_spans && !hinted; si++) {
          if (spans[si].dst_offset != off) break;
          bool found = false;
          const uint64_t* q = (const uint64_t*)(dst_base + off);
          for (uint64_t k = 0; k < spans[si].n_blocks; k++) if (!q[2 * k] && !q[2 * k + 1]) { found = true; break; }
          if (found) { which = si; break; }
          off += spans[si].n_blocks * 16;
        }
        *bad_span = which;
      }
      return fail(B200_E_BAD_BLOCK, "invalid UASTC block in input");
    }
  }
  return B200_OK;
}

int b200_uastc_encode_blocks_dev(const void* d_block_pixels, void* d_out, uint64_t n_blocks, uint32_t pack_flags, void* stream) {
  if (n_blocks && (!d_block_pixels || !d_out)) return fail(B200_E_INVALID_VALUE, "null device pointer");
  int rc = b200_uastc_init();
  if (rc) return rc;
  if (!n_blocks) return B200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (!st) { cudaStream_t* ss; rc = get_streams(&ss); if (rc) return rc; st = ss[0]; }
  const bool full = (pack_flags & 7u) >= 4u;
  if (full) k_encode_blocks<MAXC_FULL><<<encode_grid(n_blocks), encode_threads(n_blocks), 0, st>>>((const uint4*)d_block_pixels, (uint4*)d_out, n_blocks, pack_flags);
  else k_encode_blocks<MAXC_FAST><<<encode_grid(n_blocks), encode_threads(n_blocks), 0, st>>>((const uint4*)d_block_pixels, (uint4*)d_out, n_blocks, pack_flags);
  g_launches++;
  CU(cudaGetLastError());
  if (!stream) CU(cudaStreamSynchronize(st));
  return B200_OK;
}

int b200_uastc_encode_image_dev(const void* d_rgba, uint32_t width, uint32_t height, uint32_t row_pitch_bytes, void* d_out,
                                uint32_t pack_flags, void* stream, uint32_t* d_any_alpha) {
  if (!d_rgba || !d_out || !width || !height) return fail(B200_E_INVALID_VALUE, "bad image");
  const uint32_t pitch = row_pitch_bytes ? row_pitch_bytes : width * 4;
  if (pitch < width * 4) return fail(B200_E_INVALID_VALUE, "row pitch smaller than a row");
  int rc = b200_uastc_init();
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (!st) { cudaStream_t* ss; rc = get_streams(&ss); if (rc) return rc; st = ss[0]; }
  const uint32_t bx = (width + 3) / 4, by = (height + 3) / 4;
  const uint64_t n = (uint64_t)bx * by;
  if ((pack_flags & 7u) >= 4u)
    k_encode_image<MAXC_FULL><<<encode_grid(n), encode_threads(n), 0, st>>>((const uint8_t*)d_rgba, width, height, pitch, (uint4*)d_out, bx, by, pack_flags, d_any_alpha);
  else
    k_encode_image<MAXC_FAST><<<encode_grid(n), encode_threads(n), 0, st>>>((const uint8_t*)d_rgba, width, height, pitch, (uint4*)d_out, bx, by, pack_flags, d_any_alpha);
  g_launches++;
  CU(cudaGetLastError());
  if (!stream) CU(cudaStreamSynchronize(st));
  return B200_OK;
}

int b200_uastc_encode_images_dev(const b200_dev_image* images, uint32_t n_images, uint32_t pack_flags, void* stream, uint32_t* d_any_alpha) {
  if (n_images && !images) return fail(B200_E_INVALID_VALUE, "null pointer");
  int rc = b200_uastc_init();
  if (rc) return rc;
  if (!n_images) return B200_OK;
  std::vector<EncImageDesc> descs(n_images);
  uint64_t total = 0;
  for (uint32_t i = 0; i < n_images; i++) {
    const b200_dev_image& im = images[i];
    if (!im.d_rgba || !im.d_out || !im.width || !im.height) return fail(B200_E_INVALID_VALUE, "bad image");
    const uint32_t pitch = im.row_pitch_bytes ? im.row_pitch_bytes : im.width * 4;
    if (pitch < im.width * 4 || (pitch & 3u) || ((uintptr_t)im.d_rgba & 3u)) return fail(B200_E_INVALID_VALUE, "row pitch / base must be 4-byte aligned and hold a row");
    descs[i] = {(const uint8_t*)im.d_rgba, (uint4*)im.d_out, im.width, im.height, pitch, (im.width + 3) / 4, total};
    total += (uint64_t)((im.width + 3) / 4) * ((im.height + 3) / 4);
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (!st) { cudaStream_t* ss; rc = get_streams(&ss); if (rc) return rc; st = ss[0]; }
  EncImageDesc* d_descs = nullptr;
  CU(cudaMallocAsync((void**)&d_descs, n_images * sizeof(EncImageDesc), st));
  cudaError_t e = cudaMemcpyAsync(d_descs, descs.data(), n_images * sizeof(EncImageDesc), cudaMemcpyHostToDevice, st);  // pageable source: staged before return
  if (e == cudaSuccess) {
    if ((pack_flags & 7u) >= 4u) k_encode_images<MAXC_FULL><<<encode_grid(total), encode_threads(total), 0, st>>>(d_descs, n_images, total, pack_flags, d_any_alpha);
    else k_encode_images<MAXC_FAST><<<encode_grid(total), encode_threads(total), 0, st>>>(d_descs, n_images, total, pack_flags, d_any_alpha);
    g_launches++;
    e = cudaGetLastError();
  }
  cudaFreeAsync(d_descs, st);
  if (e != cudaSuccess) return fail(B200_E_CUDA, "encode_images_dev", e);
  if (!stream) CU(cudaStreamSynchronize(st));
  return B200_OK;
}

int b200_uastc_rdo_blocks_dev(void* d_blocks, const void* d_block_pixels, uint64_t n_blocks, const b200_uastc_params* params, void* stream) {
  if (!params || params->struct_size != sizeof(b200_uastc_params)) return fail(B200_E_INVALID_VALUE, "params->struct_size");
  if (n_blocks && (!d_blocks || !d_block_pixels)) return fail(B200_E_INVALID_VALUE, "null device pointer");
  int rc = b200_uastc_init();
  if (rc) return rc;
  if (!n_blocks) return B200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (!st) { cudaStream_t* ss; rc = get_streams(&ss); if (rc) return rc; st = ss[0]; }
  std::vector<uint64_t> nb(1, n_blocks);
  return rdo_run((uint4*)d_blocks, (const uint4*)d_block_pixels, nb, *params, st);
}

int b200_uastc_encode(const b200_image* images, uint32_t n_images, const b200_uastc_params* params, uint8_t* const* out_blocks,
                      uint32_t* out_any_alpha) {
  return b200_uastc_encode_ex(images, n_images, nullptr, params, out_blocks, out_any_alpha);
}

int b200_uastc_encode_ex(const b200_image* images, uint32_t n_images, const b200_src_format* format, const b200_uastc_params* params,
                         uint8_t* const* out_blocks, uint32_t* out_any_alpha) {
  if (out_any_alpha) *out_any_alpha = 0;
  SrcFmt fmt;
  if (format) {
    if (format->num_components < 1 || format->num_components > 4) return fail(B200_E_INVALID_VALUE, "num_components must be 1..4");
    fmt.comps = format->num_components; fmt.sel = 0;
    for (int c = 0; c < 4; c++) {
      const uint32_t sc = format->swizzle[c];
      if (sc < 1 || sc > 6) return fail(B200_E_INVALID_VALUE, "swizzle entries are swizzle_e values 1..6");
      if (sc <= 4 && sc > fmt.comps) return fail(B200_E_INVALID_VALUE, "swizzle names a component the source does not have");
      fmt.sel |= sc << (8 * c);
    }
  }
  if (!params || params->struct_size != sizeof(b200_uastc_params)) return fail(B200_E_INVALID_VALUE, "params->struct_size");
  if (n_images && (!images || !out_blocks)) return fail(B200_E_INVALID_VALUE, "null pointer");
  for (uint32_t i = 0; i < n_images; i++)
    if (!images[i].rgba || !images[i].width || !images[i].height || !out_blocks[i]) return fail(B200_E_INVALID_VALUE, "bad image");
  int rc = b200_uastc_init();
  if (rc) return rc;
  cudaStream_t* ss;
  rc = get_streams(&ss);
  if (rc) return rc;
  if (!n_images) return B200_OK;
  if (params->rdo_enable) return encode_with_rdo(images, n_images, fmt, params, out_blocks, out_any_alpha, ss);

  // Slab list: every image is cut into runs of whole block rows (contiguous in the RGBA source and in the UASTC
  // output) of at most SLAB_BYTES of input, so upload, encode and download of neighbouring slabs overlap on three
  // streams through a ring of staging buffers.
  const size_t SLAB_BYTES = 32u << 20;
  struct Slab { uint32_t img, y0, rows_px; size_t in_bytes, out_off, out_bytes; };
  std::vector<Slab> slabs;
  size_t max_in = 0, max_out = 0, max_exp = 0;
  for (uint32_t i = 0; i < n_images; i++) {
    const b200_image& im = images[i];
    const size_t pitch = im.row_pitch_bytes ? im.row_pitch_bytes : (size_t)im.width * fmt.comps;
    if (pitch < (size_t)im.width * fmt.comps) return fail(B200_E_INVALID_VALUE, "row pitch smaller than a row");
    const uint32_t bx = (im.width + 3) / 4, by = (im.height + 3) / 4;
    uint32_t rows_per = (uint32_t)std::max<size_t>(1, SLAB_BYTES / (std::max(pitch, (size_t)im.width * 4) * 4));
    for (uint32_t r0 = 0; r0 < by; r0 += rows_per) {
      const uint32_t r1 = std::min(by, r0 + rows_per);
      const uint32_t y0 = r0 * 4, y1 = std::min(im.height, r1 * 4);
      Slab sl{i, y0, y1 - y0, pitch * (y1 - y0), (size_t)r0 * bx * 16, (size_t)(r1 - r0) * bx * 16};
      max_in = std::max(max_in, sl.in_bytes);
      max_out = std::max(max_out, sl.out_bytes);
      max_exp = std::max(max_exp, (size_t)im.width * 4 * (y1 - y0));
      slabs.push_back(sl);
    }
  }
  const int SLOTS = 3;
  uint8_t* d_in[SLOTS] = {0}; uint8_t* d_out[SLOTS] = {0}; uint8_t* d_exp[SLOTS] = {0};
  unsigned int* d_alpha = nullptr;
  cudaEvent_t ev_in[SLOTS], ev_k[SLOTS], ev_out[SLOTS];
  const int nslots = (int)std::min<size_t>(SLOTS, slabs.size());
  auto cleanup = [&]() {};   // staging buffers stay cached (t_stage)
  for (int i = 0; i < nslots; i++) {
    d_in[i] = (uint8_t*)t_stage.get(i, max_in); d_out[i] = (uint8_t*)t_stage.get(3 + i, max_out);
    if (!fmt.identity()) d_exp[i] = (uint8_t*)t_stage.get(7 + i, max_exp);
    if (!d_in[i] || !d_out[i] || (!fmt.identity() && !d_exp[i])) return fail(B200_E_OUT_OF_MEMORY, "cudaMalloc staging");
    cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ev_k[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&ev_out[i], cudaEventDisableTiming);
  }
  d_alpha = (unsigned int*)t_stage.get(6, sizeof(unsigned int));
  if (!d_alpha) return fail(B200_E_OUT_OF_MEMORY, "cudaMalloc flag");
  // Kernels of consecutive slabs go to two streams in turn: the tail of one slab's grid (fewer CTAs left than SMs)
  // overlaps the head of the next instead of idling SMs.
  cudaStream_t s_in = ss[0], s_kk[2] = {ss[1], ss[3]}, s_out = ss[2];
  cudaEvent_t ev_zero;
  cudaEventCreateWithFlags(&ev_zero, cudaEventDisableTiming);
  cudaMemsetAsync(d_alpha, 0, sizeof(unsigned int), s_kk[0]);
  cudaEventRecord(ev_zero, s_kk[0]);
  cudaStreamWaitEvent(s_kk[1], ev_zero, 0);
  const bool full = (params->pack_flags & 7u) >= 4u;
  cudaError_t err = cudaSuccess;
  for (size_t k = 0; k < slabs.size() && err == cudaSuccess; k++) {
    const Slab& sl = slabs[k];
    const b200_image& im = images[sl.img];
    const int b = (int)(k % nslots);
    cudaStream_t s_k = s_kk[k & 1];
    uint32_t pitch = im.row_pitch_bytes ? im.row_pitch_bytes : im.width * fmt.comps;
    const uint32_t bx = (im.width + 3) / 4, by = (sl.rows_px + 3) / 4;
    if (k >= (size_t)nslots) { cudaStreamWaitEvent(s_in, ev_k[b], 0); cudaStreamWaitEvent(s_k, ev_out[b], 0); }
    err = cudaMemcpyAsync(d_in[b], im.rgba + (size_t)sl.y0 * pitch, sl.in_bytes, cudaMemcpyHostToDevice, s_in);
    if (err != cudaSuccess) break;
    cudaEventRecord(ev_in[b], s_in);
    cudaStreamWaitEvent(s_k, ev_in[b], 0);
    const uint64_t n = (uint64_t)bx * by;
    const uint8_t* d_rgba = d_in[b];
    if (!fmt.identity()) {   // component expansion / swizzle on the device: the upload carries the texture's own 1-4 bytes per texel
      k_expand_swizzle<<<expand_grid((uint64_t)im.width * sl.rows_px), 256, 0, s_k>>>(d_in[b], im.width, sl.rows_px, pitch, fmt.comps, fmt.sel, (uint32_t*)d_exp[b]);
      g_launches++;
      d_rgba = d_exp[b]; pitch = im.width * 4;
    }
    if (full) k_encode_image<MAXC_FULL><<<encode_grid(n), encode_threads(n), 0, s_k>>>(d_rgba, im.width, sl.rows_px, pitch, (uint4*)d_out[b], bx, by, params->pack_flags, d_alpha);
    else k_encode_image<MAXC_FAST><<<encode_grid(n), encode_threads(n), 0, s_k>>>(d_rgba, im.width, sl.rows_px, pitch, (uint4*)d_out[b], bx, by, params->pack_flags, d_alpha);
    g_launches++;
    cudaEventRecord(ev_k[b], s_k);
    cudaStreamWaitEvent(s_out, ev_k[b], 0);
    err = cudaMemcpyAsync(out_blocks[sl.img] + sl.out_off, d_out[b], sl.out_bytes, cudaMemcpyDeviceToHost, s_out);
    cudaEventRecord(ev_out[b], s_out);
  }
  unsigned int h_alpha = 0;
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_in);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_kk[0]);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_kk[1]);
  cudaEventDestroy(ev_zero);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s_out);
  if (err == cudaSuccess) err = cudaGetLastError();
  if (err == cudaSuccess) err = cudaMemcpy(&h_alpha, d_alpha, sizeof(h_alpha), cudaMemcpyDeviceToHost);
  for (int i = 0; i < nslots; i++) { cudaEventDestroy(ev_in[i]); cudaEventDestroy(ev_k[i]); cudaEventDestroy(ev_out[i]); }
  cleanup();
  if (err != cudaSuccess) return fail(B200_E_CUDA, "encode pipeline", err);
  if (out_any_alpha) *out_any_alpha = h_alpha ? 1u : 0u;
  return B200_OK;
}

}  // extern "C"
