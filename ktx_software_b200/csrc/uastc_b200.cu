// uastc_b200.cu -- CUDA kernels (sm_100a) and the C ABI declared in include/ktx_b200_uastc.h.
//
// Kernels: k_transcode_tiled (reference hot loop #3, XC/basisu_transcoder.cpp:10128-10305: persistent CTAs, TMA bulk tiles,
// blocks binned by UASTC mode in shared memory), k_encode_image(s) / k_encode_blocks (hot loop #1, ENC/basisu_comp.cpp:1519-1562
// -> encode_uastc), k_rdo_* (hot loop #2, ENC/basisu_uastc_enc.cpp:3825-4165), k_mip_down (2x2 box filter for the device-resident
// mip chain).  The host side (device contexts, lane pool, multi-device planner, slab pipelines that stream
// H2D -> kernel -> D2H through a ring of device buffers on three streams) is in uastc_host.inc.
#include <cuda_runtime.h>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>
#include <thread>
#include <algorithm>
#include <stdio.h>

#include "../../include/ktx_b200_uastc.h"
#include "uastc_to_bc7.cuh"
#include "uastc_to_astc.cuh"
#include "uastc_encode.cuh"
#include "uastc_rdo.cuh"
#include "uastc_to_rgba.cuh"
#include "uastc_to_bcn.cuh"
#include "uastc_jobs.cuh"

namespace {

thread_local std::string g_err;
std::atomic<uint64_t> g_launches{0};

int fail(int code, const char* what, cudaError_t e = cudaSuccess) {
  g_err = what;
  if (e != cudaSuccess) { g_err += ": "; g_err += cudaGetErrorString(e); }
  return code;
}
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(e_ == cudaErrorMemoryAllocation ? B200_E_OUT_OF_MEMORY : B200_E_CUDA, #call, e_); } while (0)

// ---------------------------------------------------------------------------------------------------------
// kernels

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
__global__ void __launch_bounds__(XT_THREADS, XT_CTAS_PER_SM) k_transcode_tiled(const uint4* __restrict__ src, uint4* __restrict__ dst, uint64_t n_blocks,
                                                                unsigned long long* __restrict__ first_bad, const XtImage im) {
  constexpr int TILE = xt_tile(TARGET), PER_THREAD = TILE / XT_THREADS;
  extern __shared__ __align__(128) unsigned char xt_smem[];
  uint4 (*s_in)[TILE] = (uint4 (*)[TILE])xt_smem;  // [2][TILE]; a block's result overwrites its own input slot
  uint4* s_out = (uint4*)(xt_smem + 2 * (size_t)TILE * 16);   // TARGET 2 only: [4 texel rows][TILE] row pieces of 16 (or 8) bytes
  uint2* s_out8 = (uint2*)(xt_smem + 2 * (size_t)TILE * 16);  // TARGET 3, 8-byte targets: the output tile
  uint64_t* s_bar = (uint64_t*)(xt_smem + xt_data_bytes(TARGET));
  constexpr bool out8 = TARGET == 3 && (KIND == 0 || KIND == 2 || KIND == 4 || KIND == 6 || KIND == 8);
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
    // Prefetch of the next tile into the other buffer: it must wait until the previous tile's bulk store has drained that
    // buffer.  Issued by thread 0 AFTER the binning (below), not here: at the top of the iteration the store is microseconds
    // old and thread 0 would stall in the wait while the other 1023 threads run into the next barrier behind it.
    auto prefetch_next = [&]() {
      if (tid == 0 && (TARGET == 3 || next < n_tiles)) {
        bulk_wait_read0();   // (TARGET 3: also frees the single 8-byte output tile before this iteration refills it)
        if (next < n_tiles) {
          const uint32_t bytes = tile_count(next) * 16u;
          mbar_expect_tx(&s_bar[buf ^ 1], bytes);
          bulk_g2s(s_in[buf ^ 1], src + next * TILE, bytes, &s_bar[buf ^ 1]);
        }
      }
    };
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
    if (TARGET == 3) prefetch_next();   // before the barrier: the 8-byte output tile must be free before anyone refills it
    __syncthreads();
    if (TARGET != 3) prefetch_next();
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
                                                      unsigned int* __restrict__ any_alpha, uint32_t vec_ok) {
  // vec_ok: base and pitch are 16-byte aligned (host-checked), so whole blocks load as four 16-byte rows
  encode_phased<MAXC>((uint64_t)bx * by, flags, out, [&](uint64_t i, uint32_t* px) {
    const uint32_t x0 = (uint32_t)(i % bx) * 4, y0 = (uint32_t)(i / bx) * 4;
    if (vec_ok && x0 + 4 <= w && y0 + 4 <= h) {
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

// Up to ENC_PACK_MAX descriptors travel by value in the kernel's parameter block (a mip chain has <= 16 levels): the launch
// then needs no device allocation and no copy.  Longer lists (cube map arrays) come from a device table.
const uint32_t ENC_PACK_MAX = 32;
struct EncDescPack { EncImageDesc d[ENC_PACK_MAX]; };

template <int MAXC, bool PACKED>
__global__ void __launch_bounds__(ENC_THREADS, ENC_MIN_CTAS) k_encode_images(const __grid_constant__ EncDescPack pack, const EncImageDesc* __restrict__ d_descs,
                                                                             uint32_t n_images, uint64_t total_blocks,
                                                                             uint32_t flags, unsigned int* __restrict__ any_alpha) {
  const EncImageDesc* descs = PACKED ? pack.d : d_descs;
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
// Job-based encode, levels 0-3 (uastc_jobs.cuh).  A CTA of EJ_T threads owns tiles of EJ_T blocks and walks each tile
// through seven stages, a CTA barrier between them:
//   load      thread = block: texels -> shared memory, block class, compacted class lists
//   estimate  warp-jobs over (estimate slot, 32 blocks of the slot's class group): partition patterns -> shared memory
//   cells     warp-jobs over (cell slot, 32 blocks): one colour-cell fit per lane -> 24 B per cell in the CTA's scratch
//   score     warp-jobs over (candidate slot, 32 blocks): assemble from the cells, score -> 8 B per candidate in the scratch
//   head      thread = block: the reference's selection rules over the stored scores, the winner decoded, BC1 / EAC hints, pack
//   etc1      warp-jobs over ((flip, colour mode), 32 blocks ordered by the cost of their search): the ETC1 hint search
//   tail      thread = block: the hint jobs' results into the packed block, store
// In the four job stages a warp claims its next warp-job from a shared counter, so every warp of the CTA is busy until
// the stage's jobs run out (the block-at-a-time kernel waited at every stage barrier for the warp that held the
// most expensive blocks), the 32 lanes of a warp-job run the SAME mode on 32 different blocks (a block-at-a-time warp that
// holds opaque and alpha blocks runs both mode lists one after the other, each under a partial mask), and a thread holds one
// cell or one candidate at a time instead of the block's whole candidate list.  The scratch (cells + scores + decoded winners + ETC1 hint job results, 1208 B per block
// of the tile) is written and read by the same CTA within one tile and reused for the next: it lives in L2.
#ifndef EJ_THREADS
#define EJ_THREADS 768
#endif
// The CTA's scratch is read and written with ld.cg / st.cg (L2 only; measured neutral against cached accesses, kept because the
// data is streamed and L1 is what holds the jobs' local frames).
template <class T> __device__ __forceinline__ T ej_ld(const T* p) { return __ldcg(p); }
template <class T> __device__ __forceinline__ void ej_st(T* p, T v) { __stcg(p, v); }
#ifndef EJ_PX_SHARED
#define EJ_PX_SHARED 1   // jobs read the block's texels in place; copying them into the local frame first measured 3-4 % slower
#endif
#ifndef EJ_PARTS_SHARED
#define EJ_PARTS_SHARED 1   // ... and the pattern list, and the block-head stage its texels: another 1 %
#endif
#ifndef EJ_MIN_CTAS
#define EJ_MIN_CTAS 1
#endif
constexpr int EJ_T = EJ_THREADS;
__host__ __device__ constexpr size_t ej_scratch_per_cta(bool full) {   // cells, scores, decoded winner, ETC1 hint jobs
  return (size_t)EJ_T * ((full ? ub::J4_NUM_CELL : ub::J_MAX_CELL_SLOTS) * 24 + (full ? ub::J4_NUM_CAND : ub::J_MAX_CAND_SLOTS) * 8 + 64 + 4 * 16);
}
constexpr size_t EJ_SCRATCH_PER_CTA = ej_scratch_per_cta(false);
struct EjShared {
  uint32_t px[EJ_T * 17];        // texel j of tile block b at [b * 17 + j]: conflict-free whether a warp reads one block or 32
  uint8_t parts[EJ_T * ub::J_PARTS];
  uint16_t list[EJ_T];           // tile blocks by class: [opaque][LA opaque][LA alpha][alpha] -> every group is one range
  uint8_t cls[EJ_T];
  uint32_t cnt[4];
  uint32_t next[4];              // next unclaimed warp-job of the estimate / cell / score / ETC1 hint stage
  uint32_t e1cnt[16];            // blocks per ETC1 cost class
  unsigned long long tile;       // the tile this CTA works on
};
// (sizeof(EjShared) + the 1 KB the system reserves per CTA must stay <= 64 KB: the next shared-memory carve-out is 100 KB, and
//  the 36 KB come out of the L1 that caches the jobs' local frames -- measured, a 40-byte overshoot cost 6 % of the kernel)
static_assert(sizeof(EjShared) + 1024 <= 65536, "EjShared crosses the 64 KB carve-out");
constexpr size_t EJ_SCRATCH_HEADER = 256;

// Stage bookkeeping of one warp: how many warp-jobs every slot has (lane l holds slots l and l + 32), as inclusive prefix sums.
struct EjStage {
  uint32_t c_lo, c_hi, incl_lo, incl_hi, total;
  __device__ __forceinline__ void build(uint32_t lane, uint32_t my_lo, uint32_t my_hi) {
    c_lo = my_lo; c_hi = my_hi;
    uint32_t v = my_lo;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, v, o); if ((int)lane >= o) v += t; }
    incl_lo = v;
    const uint32_t tot_lo = __shfl_sync(0xFFFFFFFFu, v, 31);
    v = my_hi;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, v, o); if ((int)lane >= o) v += t; }
    incl_hi = tot_lo + v;
    total = __shfl_sync(0xFFFFFFFFu, incl_hi, 31);
  }
  // warp-job w -> its slot and the index of its first block inside the slot's group
  __device__ __forceinline__ void decode(uint32_t w, uint32_t& slot, uint32_t& r0) const {
    const uint32_t m_lo = __ballot_sync(0xFFFFFFFFu, w < incl_lo);
    uint32_t excl;
    if (m_lo) { slot = __ffs(m_lo) - 1; excl = __shfl_sync(0xFFFFFFFFu, incl_lo - c_lo, slot); }
    else { const uint32_t m_hi = __ballot_sync(0xFFFFFFFFu, w < incl_hi); const uint32_t l = __ffs(m_hi) - 1; slot = 32 + l; excl = __shfl_sync(0xFFFFFFFFu, incl_hi - c_hi, l); }
    r0 = (w - excl) * 32;
  }
};

// Stage barriers of k_encode_jobs.  A build with -DEJ_STAGE_CLOCKS=1 (development only) adds up, per barrier, the cycles lane 0 of
// every warp spent before arriving and until released: busy / (warps x wall) is what the stage's tail costs (tools/stage_clocks.py).
#ifndef EJ_STAGE_CLOCKS
#define EJ_STAGE_CLOCKS 0
#endif
#if EJ_STAGE_CLOCKS
__device__ unsigned long long g_ej_clk[2][10];
#define EJ_SYNC(k) do { const long long t_a_ = clock64(); __syncthreads(); const long long t_b_ = clock64(); \
    if (lane == 0) { ej_busy[k] += (unsigned long long)(t_a_ - ej_t0); ej_wall[k] += (unsigned long long)(t_b_ - ej_t0); } ej_t0 = t_b_; } while (0)
#else
#define EJ_SYNC(k) __syncthreads()
#endif

// FULL = level 4 (no partition estimates: the j4_* slot enumeration of uastc_jobs.cuh, 343 cell and 170 candidate slots per block).
template <bool PACKED, bool FULL = false>
__global__ void __launch_bounds__(EJ_THREADS, EJ_MIN_CTAS) k_encode_jobs(const __grid_constant__ EncDescPack pack, const EncImageDesc* __restrict__ d_descs,
                                                                         uint32_t n_images, uint64_t total_blocks, uint32_t flags,
                                                                         unsigned int* __restrict__ any_alpha, uint2* scratch_all) {
  extern __shared__ __align__(16) unsigned char ej_smem[];
  EjShared& sh = *(EjShared*)ej_smem;
  const EncImageDesc* descs = PACKED ? pack.d : d_descs;
  const uint32_t tid = threadIdx.x, lane = tid & 31u;
  unsigned long long* tile_counter = (unsigned long long*)scratch_all;          // zeroed by the host before the launch
  constexpr size_t EJ_SCRATCH_BYTES = ej_scratch_per_cta(FULL);
  constexpr uint32_t CELL_SLOTS = FULL ? ub::J4_NUM_CELL : ub::J_MAX_CELL_SLOTS, CAND_SLOTS = FULL ? ub::J4_NUM_CAND : ub::J_MAX_CAND_SLOTS;
  uint2* cells = scratch_all + EJ_SCRATCH_HEADER / 8 + (size_t)blockIdx.x * (EJ_SCRATCH_BYTES / 8);   // [cell slot][3][EJ_T]
  uint2* scores = cells + (size_t)CELL_SLOTS * 3 * EJ_T;                          // [candidate slot][EJ_T]
  uint32_t* gdec = (uint32_t*)(scores + (size_t)CAND_SLOTS * EJ_T);               // [texel][EJ_T]: every block's winner, decoded
  uint4* ge1 = (uint4*)(gdec + 16 * EJ_T);                                        // [ETC1 hint job slot][EJ_T]: {err lo, err hi, hint, -}
  // (cells / scores / gdec / ge1 live in global memory, i.e. in L2, not in shared memory: every KB of shared memory comes out of
  //  the L1 that holds the jobs' local frames; with gdec and ge1 in shared memory the kernel ran 19 % slower)
  ub::EncCfg cfg = ub::enc_cfg_from_flags(flags);
  // enabled slots (uniform over the launch)
  const uint32_t en_cands = ub::j_enabled_cands(cfg.mode_mask);
  uint64_t en_cells = 0;
  for (uint32_t gc = 0; gc < ub::J_NUM_CELL; gc++) if ((en_cands >> (ub::UBT(j_cell)[gc] & 31u)) & 1u) en_cells |= 1ull << gc;
  uint32_t en_est = 0;
  for (uint32_t e = 0; e < ub::J_NUM_EST; e++) if ((cfg.mode_mask >> (ub::UBT(j_est)[e] & 31u)) & 1u) en_est |= 1u << e;

  const uint64_t n_tiles = (total_blocks + EJ_T - 1) / EJ_T;
#if EJ_STAGE_CLOCKS
  unsigned long long ej_busy[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, ej_wall[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  long long ej_t0 = clock64();
#endif
  for (;;) {
    // tiles are claimed from a global counter: tile costs differ several-fold (solid / opaque / alpha content), a fixed share
    // per CTA leaves SMs idle behind the CTA that drew the expensive tiles
    if (tid == 0) sh.tile = atomicAdd(tile_counter, 1ull);
    if (tid < 4) sh.cnt[tid] = 0;
    if (tid < 4) sh.next[tid] = 0;
    if (tid < 16) sh.e1cnt[tid] = 0;
    EJ_SYNC(0);
    const uint64_t tile = sh.tile;
#if EJ_STAGE_CLOCKS
    if (tile >= n_tiles) { if (lane == 0) for (int k = 0; k < 10; k++) { atomicAdd(&g_ej_clk[0][k], ej_busy[k]); atomicAdd(&g_ej_clk[1][k], ej_wall[k]); } }
#endif
    if (tile >= n_tiles) break;   // uniform per CTA
    const uint64_t base = tile * EJ_T;
    // ---- load: thread = block
    const uint64_t i = base + tid;
    const bool live = i < total_blocks;
    uint4* dst = nullptr;
    uint32_t cls = 0;
    if (live) {
      uint32_t lo = 0, hi = n_images - 1;  // last image whose first_block <= i
      while (lo < hi) { const uint32_t mid = (lo + hi + 1) >> 1; if (descs[mid].first_block <= i) lo = mid; else hi = mid - 1; }
      const EncImageDesc d = descs[lo];
      uint32_t px[16];
      uint32_t amin = 255;
      if (d.w == 0) {   // pre-gathered 64-byte texel blocks (b200_uastc_encode_blocks_dev)
        const uint64_t k = i - d.first_block;
        dst = d.out + k;
#pragma unroll
        for (uint32_t y = 0; y < 4; y++) {
          const uint4 v = __ldg((const uint4*)d.img + k * 4 + y);
          px[4 * y] = v.x; px[4 * y + 1] = v.y; px[4 * y + 2] = v.z; px[4 * y + 3] = v.w;
        }
      } else {
        const uint64_t k = block_of_thread(i - d.first_block, d.bx, (d.h + 3) / 4);
        dst = d.out + k;
        const uint32_t x0 = (uint32_t)(k % d.bx) * 4, y0 = (uint32_t)(k / d.bx) * 4;
        if (x0 + 4 <= d.w && y0 + 4 <= d.h && !(d.pitch & 15u) && !((uintptr_t)d.img & 15u)) {
#pragma unroll
          for (uint32_t y = 0; y < 4; y++) {
            const uint4 v = __ldg((const uint4*)(d.img + (size_t)(y0 + y) * d.pitch + (size_t)x0 * 4));
            px[4 * y] = v.x; px[4 * y + 1] = v.y; px[4 * y + 2] = v.z; px[4 * y + 3] = v.w;
          }
        } else {
          for (uint32_t y = 0; y < 4; y++)
            for (uint32_t x = 0; x < 4; x++) {
              const uint32_t sx = min(x0 + x, d.w - 1), sy = min(y0 + y, d.h - 1);
              px[4 * y + x] = __ldg((const uint32_t*)(d.img + (size_t)sy * d.pitch + (size_t)sx * 4));
            }
        }
      }
#pragma unroll
      for (uint32_t j = 0; j < 16; j++) { sh.px[tid * 17 + j] = px[j]; amin = min(amin, px[j] >> 24); }
      if (amin < 255 && any_alpha) atomicOr(any_alpha, 1u);
      cls = ub::j_classify(px, cfg);
    }
    sh.cls[tid] = (uint8_t)cls;
    // class lists: rank inside the class from a warp-aggregated shared atomic
    const uint32_t ci = !(cls & ub::J_CLS_LIVE) ? 4u : ((cls & ub::J_CLS_LA) ? ((cls & ub::J_CLS_ALPHA) ? 2u : 1u) : ((cls & ub::J_CLS_ALPHA) ? 3u : 0u));
    uint32_t my_rank = 0;
    {
      const uint32_t peers = __match_any_sync(0xFFFFFFFFu, ci);
      const uint32_t leader = __ffs(peers) - 1;
      uint32_t b0 = 0;
      if (lane == leader && ci < 4u) b0 = atomicAdd(&sh.cnt[ci], __popc(peers));
      b0 = __shfl_sync(0xFFFFFFFFu, b0, leader);
      my_rank = b0 + __popc(peers & ((1u << lane) - 1u));
    }
    EJ_SYNC(1);
    const uint32_t c0 = sh.cnt[0], c1 = sh.cnt[1], c2 = sh.cnt[2], c3 = sh.cnt[3];
    if (ci < 4u) sh.list[(ci == 0 ? 0u : ci == 1 ? c0 : ci == 2 ? c0 + c1 : c0 + c1 + c2) + my_rank] = (uint16_t)tid;
    // group -> range of the list: LA = [c0, c0 + c1 + c2), opaque = [0, c0 + c1), alpha = [c0 + c1, all)
    // (level 4 tries the alpha modes on every live block: its "alpha" group is the whole list)
    const uint32_t g_start[3] = {c0, 0u, FULL ? 0u : c0 + c1}, g_count[3] = {c1 + c2, c0 + c1, FULL ? c0 + c1 + c2 + c3 : c2 + c3};
    auto warps_of = [&](uint32_t g) { return (g_count[g] + 31u) >> 5; };
    EJ_SYNC(2);

    // ---- estimate jobs
    if (!FULL) {
      EjStage stg;
      stg.build(lane, (lane < ub::J_NUM_EST && ((en_est >> lane) & 1u)) ? warps_of(ub::UBT(j_est)[lane] >> 22) : 0u, 0u);
      for (;;) {
        uint32_t w = 0;
        if (lane == 0) w = atomicAdd(&sh.next[0], 1u);
        w = __shfl_sync(0xFFFFFFFFu, w, 0);
        if (w >= stg.total) break;
        uint32_t e, r0;
        stg.decode(w, e, r0);
        const uint32_t g = ub::UBT(j_est)[e] >> 22, r = r0 + lane;
        if (r < g_count[g]) {
          const uint32_t b = sh.list[g_start[g] + r];
#if EJ_PX_SHARED
          const uint32_t* px = sh.px + b * 17;   // read in place (generic loads from shared memory): no copy into the local frame
#else
          uint32_t px[16];
#pragma unroll
          for (uint32_t j = 0; j < 16; j++) px[j] = sh.px[b * 17 + j];
#endif
          ub::j_run_estimate(e, px, sh.parts + b * ub::J_PARTS);
        }
      }
    }
    EJ_SYNC(3);

    // ---- cell jobs
    if (FULL) {
      // warp-job = (mode entry, candidate, cell, 32 blocks of the entry's group): lane e < 18 holds entry e's job count
      EjStage stg;
      stg.build(lane, lane < ub::J4_NUM_MODES ? ub::j4_count(lane) * ub::j4_cells(lane) * warps_of(ub::j4_group(lane)) : 0u, 0u);
      for (;;) {
        uint32_t w = 0;
        if (lane == 0) w = atomicAdd(&sh.next[1], 1u);
        w = __shfl_sync(0xFFFFFFFFu, w, 0);
        if (w >= stg.total) break;
        uint32_t e, r0;
        stg.decode(w, e, r0);
        const uint32_t g = ub::j4_group(e), chunks = warps_of(g), local = r0 >> 5, slot = local / chunks, r = (local - slot * chunks) * 32 + lane;
        const uint32_t k = slot / ub::j4_cells(e), cell = slot - k * ub::j4_cells(e);
        if (r < g_count[g]) {
          const uint32_t b = sh.list[g_start[g] + r];
          ub::CellOut o;
          ub::j4_run_cell(e, k, cell, sh.px + b * 17, cfg, o);
          uint2* q = cells + (size_t)(ub::UBT(j4_cell_base)[e] + slot) * 3 * EJ_T + b;
          ej_st(q, make_uint2(o.astc_lo, o.astc_hi));
          ej_st(q + EJ_T, ((const uint2*)o.sel)[0]);
          ej_st(q + 2 * EJ_T, ((const uint2*)o.sel)[1]);
        }
      }
    } else {
      EjStage stg;
      const uint32_t s_hi = lane + 32u;
      stg.build(lane, ((en_cells >> lane) & 1ull) ? warps_of(ub::jc_group(ub::UBT(j_cand)[ub::UBT(j_cell)[lane] & 31u])) : 0u,
                (s_hi < ub::J_NUM_CELL && ((en_cells >> s_hi) & 1ull)) ? warps_of(ub::jc_group(ub::UBT(j_cand)[ub::UBT(j_cell)[s_hi] & 31u])) : 0u);
      for (;;) {
        uint32_t w = 0;
        if (lane == 0) w = atomicAdd(&sh.next[1], 1u);
        w = __shfl_sync(0xFFFFFFFFu, w, 0);
        if (w >= stg.total) break;
        uint32_t gc, r0;
        stg.decode(w, gc, r0);
        const uint32_t g = ub::jc_group(ub::UBT(j_cand)[ub::UBT(j_cell)[gc] & 31u]), r = r0 + lane;
        if (r < g_count[g]) {
          const uint32_t b = sh.list[g_start[g] + r];
#if EJ_PX_SHARED
          const uint32_t* px = sh.px + b * 17;   // read in place (generic loads from shared memory): no copy into the local frame
#else
          uint32_t px[16];
#pragma unroll
          for (uint32_t j = 0; j < 16; j++) px[j] = sh.px[b * 17 + j];
#endif
#if EJ_PARTS_SHARED
          const uint8_t* parts = sh.parts + b * ub::J_PARTS;
#else
          alignas(4) uint8_t parts[ub::J_PARTS];
#pragma unroll
          for (uint32_t j = 0; j < ub::J_PARTS / 4; j++) ((uint32_t*)parts)[j] = ((const uint32_t*)(sh.parts + b * ub::J_PARTS))[j];
#endif
          ub::CellOut o;
          ub::j_run_cell(gc, px, parts, cfg, o);
          uint2* q = cells + (size_t)ub::j_cell_local(gc) * 3 * EJ_T + b;
          ej_st(q, make_uint2(o.astc_lo, o.astc_hi));
          ej_st(q + EJ_T, ((const uint2*)o.sel)[0]);
          ej_st(q + 2 * EJ_T, ((const uint2*)o.sel)[1]);
        }
      }
    }
    EJ_SYNC(4);   // (global scratch written above is read below by other threads of this CTA: the barrier orders it)

    // ---- candidate jobs: assemble + score
    if (FULL) {
      EjStage stg;
      stg.build(lane, lane < ub::J4_NUM_MODES ? ub::j4_count(lane) * warps_of(ub::j4_group(lane)) : 0u, 0u);
      for (;;) {
        uint32_t w = 0;
        if (lane == 0) w = atomicAdd(&sh.next[2], 1u);
        w = __shfl_sync(0xFFFFFFFFu, w, 0);
        if (w >= stg.total) break;
        uint32_t e, r0;
        stg.decode(w, e, r0);
        const uint32_t g = ub::j4_group(e), chunks = warps_of(g), local = r0 >> 5, k = local / chunks, r = (local - k * chunks) * 32 + lane;
        if (r < g_count[g]) {
          const uint32_t b = sh.list[g_start[g] + r];
          ub::CellOut co[3];
          const uint32_t nc = ub::j4_cells(e);
          for (uint32_t s = 0; s < nc; s++) {
            const uint2* q = cells + (size_t)(ub::UBT(j4_cell_base)[e] + k * nc + s) * 3 * EJ_T + b;
            const uint2 a = ej_ld(q), s0 = ej_ld(q + EJ_T), s1 = ej_ld(q + 2 * EJ_T);
            co[s].astc_lo = a.x; co[s].astc_hi = a.y;
            ((uint2*)co[s].sel)[0] = s0; ((uint2*)co[s].sel)[1] = s1;
          }
          ub::Candidate c;
          ub::j4_assemble(e, k, co, c);
          uint32_t overall;
          float uerr;
          ub::j_score(sh.px + b * 17, c, sh.cls[b], flags, overall, uerr);
          ej_st(scores + (size_t)(ub::UBT(j4_cand_base)[e] + k) * EJ_T + b, make_uint2(overall, __float_as_uint(uerr)));
        }
      }
    } else {
      EjStage stg;
      stg.build(lane, ((en_cands >> lane) & 1u) ? warps_of(ub::jc_group(ub::UBT(j_cand)[lane])) : 0u, 0u);
      for (;;) {
        uint32_t w = 0;
        if (lane == 0) w = atomicAdd(&sh.next[2], 1u);
        w = __shfl_sync(0xFFFFFFFFu, w, 0);
        if (w >= stg.total) break;
        uint32_t gs, r0;
        stg.decode(w, gs, r0);
        const uint32_t d = ub::UBT(j_cand)[gs], g = ub::jc_group(d), r = r0 + lane;
        if (r < g_count[g]) {
          const uint32_t b = sh.list[g_start[g] + r];
#if EJ_PX_SHARED
          const uint32_t* px = sh.px + b * 17;   // read in place (generic loads from shared memory): no copy into the local frame
#else
          uint32_t px[16];
#pragma unroll
          for (uint32_t j = 0; j < 16; j++) px[j] = sh.px[b * 17 + j];
#endif
#if EJ_PARTS_SHARED
          const uint8_t* parts = sh.parts + b * ub::J_PARTS;
#else
          alignas(4) uint8_t parts[ub::J_PARTS];
#pragma unroll
          for (uint32_t j = 0; j < ub::J_PARTS / 4; j++) ((uint32_t*)parts)[j] = ((const uint32_t*)(sh.parts + b * ub::J_PARTS))[j];
#endif
          ub::CellOut co[3];
          const uint32_t nc = ub::jc_cells(d);
          for (uint32_t s = 0; s < nc; s++) {
            const uint2* q = cells + (size_t)ub::j_cell_local(ub::jc_cell0(d) + s) * 3 * EJ_T + b;
            const uint2 a = ej_ld(q), s0 = ej_ld(q + EJ_T), s1 = ej_ld(q + 2 * EJ_T);
            co[s].astc_lo = a.x; co[s].astc_hi = a.y;
            ((uint2*)co[s].sel)[0] = s0; ((uint2*)co[s].sel)[1] = s1;
          }
          ub::Candidate c;
          ub::j_assemble(gs, parts, co, c);
          uint32_t overall;
          float uerr;
          ub::j_score(px, c, sh.cls[b], flags, overall, uerr);
          ej_st(scores + (size_t)ub::j_cand_local(gs) * EJ_T + b, make_uint2(overall, __float_as_uint(uerr)));
        }
      }
    }
    EJ_SYNC(5);

    // ---- finish, head: thread = block.  Selection, the winner decoded, BC1 / EAC hints, the block packed but for its ETC1 hints
    ub::Bits128 packed;
    uint32_t etc1_at;
    {
      const bool solid = cls == ub::J_CLS_SOLID;
      ub::Candidate win;
#if EJ_PARTS_SHARED
      const uint32_t* px = sh.px + tid * 17;
#else
      uint32_t px[16];
#pragma unroll
      for (uint32_t j = 0; j < 16; j++) px[j] = sh.px[tid * 17 + j];
#endif
      if (FULL) {
        if (live && !solid) {
          // the block's candidates in list order: the entries of its groups, every candidate of each
          uint32_t overall[ub::J4_NUM_CAND], n = 0;
          float uerr[ub::J4_NUM_CAND];
          uint8_t modes[ub::J4_NUM_CAND], slots[ub::J4_NUM_CAND];
          for (uint32_t e = 0; e < ub::J4_NUM_MODES; e++) {
            if (!ub::j4_block_in_group(cls, ub::j4_group(e))) continue;
            const uint32_t c0s = ub::UBT(j4_cand_base)[e], cnt = ub::j4_count(e);
            for (uint32_t k = 0; k < cnt; k++) {
              const uint2 v = ej_ld(scores + (size_t)(c0s + k) * EJ_T + tid);
              overall[n] = v.x; uerr[n] = __uint_as_float(v.y);
              modes[n] = (uint8_t)ub::j4_mode(e); slots[n] = (uint8_t)(c0s + k); n++;
            }
          }
          const uint32_t gs = slots[ub::j_select(n, overall, uerr, modes, flags)];
          uint32_t e = 0;
          while (ub::UBT(j4_cand_base)[e + 1] <= gs) e++;
          const uint32_t k = gs - ub::UBT(j4_cand_base)[e], nc = ub::j4_cells(e);
          ub::CellOut co[3];
          for (uint32_t s = 0; s < nc; s++) {
            const uint2* q = cells + (size_t)(ub::UBT(j4_cell_base)[e] + k * nc + s) * 3 * EJ_T + tid;
            const uint2 a = ej_ld(q), s0 = ej_ld(q + EJ_T), s1 = ej_ld(q + 2 * EJ_T);
            co[s].astc_lo = a.x; co[s].astc_hi = a.y;
            ((uint2*)co[s].sel)[0] = s0; ((uint2*)co[s].sel)[1] = s1;
          }
          ub::j4_assemble(e, k, co, win);
        }
      } else if (live && !solid) {
        const uint32_t mine = ub::j_block_cands(cls, en_cands);
        uint32_t overall[ub::J_MAX_CAND_SLOTS], n = 0;
        float uerr[ub::J_MAX_CAND_SLOTS];
        uint8_t modes[ub::J_MAX_CAND_SLOTS], slots[ub::J_MAX_CAND_SLOTS];
        for (uint32_t m = mine; m; m &= m - 1) {
          const uint32_t gs = __ffs(m) - 1;
          const uint2 v = ej_ld(scores + (size_t)ub::j_cand_local(gs) * EJ_T + tid);
          overall[n] = v.x; uerr[n] = __uint_as_float(v.y);
          modes[n] = (uint8_t)ub::jc_mode(ub::UBT(j_cand)[gs]); slots[n] = (uint8_t)gs; n++;
        }
        const uint32_t gs = slots[ub::j_select(n, overall, uerr, modes, flags)];
        const uint32_t d = ub::UBT(j_cand)[gs], nc = ub::jc_cells(d);
        ub::CellOut co[3];
        for (uint32_t s = 0; s < nc; s++) {
          const uint2* q = cells + (size_t)ub::j_cell_local(ub::jc_cell0(d) + s) * 3 * EJ_T + tid;
          const uint2 a = ej_ld(q), s0 = ej_ld(q + EJ_T), s1 = ej_ld(q + 2 * EJ_T);
          co[s].astc_lo = a.x; co[s].astc_hi = a.y;
          ((uint2*)co[s].sel)[0] = s0; ((uint2*)co[s].sel)[1] = s1;
        }
        ub::j_assemble(gs, sh.parts + tid * ub::J_PARTS, co, win);
      }
      uint32_t dec[16], wmode;
      packed = ub::j_finish_head(px, flags, &win, live, solid, ENC_FINE_SYNC, dec, wmode, etc1_at);
      uint32_t e1cls = 16;
      if (etc1_at != 0xFFFFu) {
#pragma unroll
        for (uint32_t j = 0; j < 16; j++) ej_st(gdec + j * EJ_T + tid, dec[j]);
        sh.parts[tid * ub::J_PARTS] = (uint8_t)wmode;   // the pattern list is not needed any more: its first byte carries the winner's mode
        e1cls = ub::etc1_cost_class(wmode, dec);
      }
      // the class lists are done with: sh.list now takes the non-solid blocks ordered by the cost class of their ETC1 search, dearest
      // first, so that the 32 lanes of a hint job run searches of similar length (same two-step ranking as the load stage)
      {
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, e1cls);
        const uint32_t leader = __ffs(peers) - 1;
        uint32_t b0 = 0;
        if (lane == leader && e1cls < 16u) b0 = atomicAdd(&sh.e1cnt[e1cls], __popc(peers));
        b0 = __shfl_sync(0xFFFFFFFFu, b0, leader);
        const uint32_t rank = b0 + __popc(peers & ((1u << lane) - 1u));
        EJ_SYNC(6);
        if (e1cls < 16u) {
          uint32_t off = 0;
          for (uint32_t k = 0; k < e1cls; k++) off += sh.e1cnt[k];
          sh.list[off + rank] = (uint16_t)tid;
        }
      }
    }
    EJ_SYNC(7);

    // ---- ETC1 hint jobs: warp-jobs over ((flip slot, colour mode), 32 non-solid blocks).  The search is the largest single item
    // of the encoder and its cost differs 20-fold between blocks (modes with and without a bias field, the range-dependent
    // table limit): as jobs, the lanes of a warp run one (flip, colour mode) pair and no warp waits for one unlucky block.
#ifndef EJ_ETC1_SPLIT_MODES
#define EJ_ETC1_SPLIT_MODES 1
#endif
    const uint32_t e1_slots = ub::etc1_job_slots(cfg.level, cfg.flags), e1_nf = e1_slots & 3u, e1_ni = EJ_ETC1_SPLIT_MODES ? e1_slots >> 2 : 1u;
    {
      const uint32_t n_blocks = c0 + c1 + c2 + c3, per_slot = (n_blocks + 31u) >> 5, total = per_slot * e1_nf * e1_ni;
      for (;;) {
        uint32_t w = 0;
        if (lane == 0) w = atomicAdd(&sh.next[3], 1u);
        w = __shfl_sync(0xFFFFFFFFu, w, 0);
        if (w >= total) break;
        const uint32_t n_slots = e1_nf * e1_ni, slot = w % n_slots, r = (w / n_slots) * 32 + lane;   // dearest chunks first, all their slots
        const uint32_t together = __ballot_sync(0xFFFFFFFFu, r < n_blocks);
        if (r < n_blocks) {
          const uint32_t b = sh.list[r];
#if EJ_PX_SHARED
          const uint32_t* px = sh.px + b * 17;
          uint32_t dec[16];
#pragma unroll
          for (uint32_t j = 0; j < 16; j++) dec[j] = ej_ld(gdec + j * EJ_T + b);
#else
          uint32_t px[16], dec[16];
#pragma unroll
          for (uint32_t j = 0; j < 16; j++) { px[j] = sh.px[b * 17 + j]; dec[j] = ej_ld(gdec + j * EJ_T + b); }
#endif
          ub::Etc1JobOut o;
          ub::etc1_hints_job(sh.parts[b * ub::J_PARTS], px, dec, cfg.level, cfg.flags, slot / e1_ni, EJ_ETC1_SPLIT_MODES ? slot % e1_ni : ub::ETC1_JOB_BOTH_MODES, o, together);
          ej_st(ge1 + slot * EJ_T + b, make_uint4((uint32_t)o.err, (uint32_t)(o.err >> 32), o.hint, 0u));
        }
      }
    }
    EJ_SYNC(8);

    // ---- finish, tail: thread = block
    if (live) {
      if (etc1_at != 0xFFFFu) {
        ub::Etc1JobOut jobs[4];
        for (uint32_t k = 0; k < e1_nf * e1_ni; k++) {
          const uint4 v = ej_ld(ge1 + k * EJ_T + tid);
          jobs[k].err = (uint64_t)v.x | ((uint64_t)v.y << 32); jobs[k].hint = v.z; jobs[k].pad_ = 0;
        }
        ub::j_finish_tail(packed, etc1_at, jobs, e1_nf * e1_ni);
      }
      *dst = make_uint4((uint32_t)packed.lo, (uint32_t)(packed.lo >> 32), (uint32_t)packed.hi, (uint32_t)(packed.hi >> 32));
    }
    EJ_SYNC(9);
  }
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
  uint32_t slot_off;         // first table slot of this chain: prefix sum of min(length, RDO_CHUNK) over the chains before it
};
const int RDO_THREADS = 320;       // 256 trial threads + the keep-cost warp (main group, 288) + the re-fit worker warp
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

// chunk-local table slot of block `bi` of chain `c`: chains[c].slot_off + (bi - first - step0)
__global__ void __launch_bounds__(256) k_rdo_prep(const uint4* __restrict__ blocks, const uint4* __restrict__ px, const RdoChain* __restrict__ chains,
                                                  uint32_t n_chains, uint32_t n_slots, uint32_t step0, ub::RdoBlockInfo* __restrict__ info,
                                                  ub::RdoPre* __restrict__ pre, uint32_t* __restrict__ tables,
                                                  unsigned int* __restrict__ err_flag, ub::RdoParams prm) {
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
  if (warp >= n_slots) return;
  uint32_t c = 0, hi = n_chains - 1;   // last chain whose slot_off <= warp
  while (c < hi) { const uint32_t mid = (c + hi + 1) >> 1; if (chains[mid].slot_off <= warp) c = mid; else hi = mid - 1; }
  const uint32_t k = warp - chains[c].slot_off;
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
  const size_t slot = (size_t)chains[c].slot_off + k;
  if (lane == 0) { info[slot] = f; if (!ok) atomicOr(err_flag, 1u); }
  if (!ok || f.mode == ub::UASTC_MODE_SOLID) return;
  uint32_t* table = tables + slot * ub::RDO_TABLE_WORDS;
  if (lane < (1u << f.wb)) ub::rdo_fill_level(p, u, f, lane, table);
  __syncwarp();  // orders the lanes' table writes before lane 0's reads
  if (lane == 0) {
    ub::RdoCur cur;
    ub::rdo_prepare_fast(p, blk, f, table, prm, cur);
    pre[slot] = ub::rdo_pre_of(cur);
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

// Threads 0..255 score the window's candidates, thread 256 prices "keep as is" (together: the 288 MAIN threads, which meet at
// a named barrier), and the last warp is the RE-FIT WORKER: the mode-0 endpoint re-fit of a rewritten block (a full forced-selector
// colour-cell fit, ~15 us on one thread) used to run on thread 0 inside the step whenever a later block might read the
// endpoint bits (6.4 of 10.2 us per step on img_mix).  It needs only the block's own texels and bits, and only a block of mode
// 15 / 17 / 18 ever reads another block's endpoint bits, so thread 0 now queues it, the worker re-fits up to 32 queued blocks at
// a time (one lane each) behind the walk, and a step whose own block is one of those reader modes first waits for the queue to
// drain -- the bits every step sees are the reference's.
const int RDO_MAIN_THREADS = 288, RDO_QUEUE = 64;
__device__ __forceinline__ void rdo_main_barrier() { asm volatile("bar.sync 1, 288;" ::: "memory"); }

__global__ void __launch_bounds__(RDO_THREADS) k_rdo_chain(uint4* blocks, const uint4* __restrict__ px, const RdoChain* __restrict__ chains,
                                                           uint32_t step0, const ub::RdoBlockInfo* __restrict__ info,
                                                           const ub::RdoPre* __restrict__ pre, const uint32_t* __restrict__ tables,
                                                           ub::RdoHistEntry* __restrict__ hist_all, const uint8_t* __restrict__ hazard,
                                                           uint8_t* __restrict__ modified, ub::RdoParams p) {
  // per-step inputs are double-buffered: while step bi is scored, the loads of step bi+1 are in flight
  __shared__ __align__(16) uint32_t s_table[2][ub::RDO_TABLE_WORDS];
  __shared__ __align__(16) uint4 s_blk[2];
  __shared__ __align__(16) ub::RdoPre s_pre[2];
  __shared__ ub::RdoBlockInfo s_info[2];
  __shared__ __align__(16) uint4 s_ring[256];
  __shared__ float s_keep_t;
  __shared__ unsigned long long s_best;
  // re-fit queue: thread 0 pushes (head), the worker warp completes (tail); entry e lives in slot e % RDO_QUEUE
  __shared__ __align__(16) uint4 s_q_blk[RDO_QUEUE];
  __shared__ uint32_t s_q_bi[RDO_QUEUE];
  __shared__ volatile uint32_t s_q_head, s_q_tail, s_q_done;

  const RdoChain ch = chains[blockIdx.x];
  const uint32_t tid = threadIdx.x;
  ub::RdoHistEntry* hist = hist_all + ch.hist_off;
  const uint32_t begin = ch.first + step0;
  if (begin >= ch.last) return;   // (uniform over the CTA)
  const uint32_t end = min(ch.last, begin + RDO_CHUNK);
  if (tid == 0) { s_q_head = 0; s_q_tail = 0; s_q_done = 0; s_best = ~0ull; }
  // refill the ring with the (final) blocks just before this chunk
  if (tid < 256) { const uint32_t back = tid + 1; if (begin >= ch.first + back) s_ring[(begin - back) & 255u] = blocks[begin - back]; }
  __syncthreads();

  if (tid >= RDO_MAIN_THREADS) {
    // ---- the re-fit worker: lane l takes queue entry tail + l
    const uint32_t lane = tid - RDO_MAIN_THREADS;
    uint32_t tail = 0;
    for (;;) {
      const uint32_t head = s_q_head;
      if (head == tail) {
        if (s_q_done) break;
        __nanosleep(200);
        continue;
      }
      __threadfence_block();   // the entries below head were written before head moved
      const uint32_t n = min(head - tail, 32u);
      if (lane < n) {
        const uint32_t slot = (tail + lane) % RDO_QUEUE, bi = s_q_bi[slot];
        uint32_t q[16];
#pragma unroll
        for (uint32_t y = 0; y < 4; y++) {
          const uint4 v = __ldg(px + (size_t)bi * 4 + y);
          q[4 * y] = v.x; q[4 * y + 1] = v.y; q[4 * y + 2] = v.z; q[4 * y + 3] = v.w;
        }
        const uint4 r = from_bits(ub::rdo_refine_mode0(q, to_bits(s_q_blk[slot])));
        blocks[bi] = r;
        s_ring[bi & 255u] = r;   // (thread 0 does not reuse the ring slot of a block whose re-fit is pending)
      }
      __syncwarp();
      __threadfence_block();
      tail += n;
      if (lane == 0) s_q_tail = tail;
    }
    return;
  }

  // what each main thread fetches for a step: 128 threads the table, one each the block, info and pre-computed stats
  uint4 nx = make_uint4(0, 0, 0, 0);
  ub::RdoBlockInfo nx_info;
  ub::RdoPre nx_pre;
  auto fetch = [&](uint32_t bi) {
    const size_t slot = (size_t)ch.slot_off + (bi - begin);
    if (tid < 128) nx = __ldg((const uint4*)(tables + slot * ub::RDO_TABLE_WORDS) + tid);
    else if (tid == 132) nx = blocks[bi];
    else if (tid == 133) nx_info = info[slot];
    else if (tid == 134) nx_pre = pre[slot];
  };
  auto stash = [&](uint32_t b) {
    if (tid < 128) ((uint4*)s_table[b])[tid] = nx;
    else if (tid == 132) s_blk[b] = nx;
    else if (tid == 133) s_info[b] = nx_info;
    else if (tid == 134) s_pre[b] = nx_pre;
  };
  fetch(begin);
  stash(0);
  rdo_main_barrier();

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
      // this block reads endpoint bits of earlier mode-0 blocks: every queued re-fit must have landed (head is stable here:
      // thread 0 pushes only between the two barriers of a step)
      if (f.valid && ub::rdo_mode_reads_mode0_endpoints(f.mode)) {
        const uint32_t head = s_q_head;
        while (s_q_tail != head) __nanosleep(100);
        __threadfence_block();
      }
      const uint32_t lo_index = (bi - ch.first > p.window) ? bi - p.window : ch.first;
      const uint32_t n_win = bi - lo_index;
      // thread k takes candidates k, k+256, ... (k = 0 is the nearest earlier block); thread 256 prices "keep as is"
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
      rdo_main_barrier();
    }
    if (tid == 0) {
      ub::Bits128 best = blk;
      bool queued = false;
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
              if (hazard[bi]) queued = true;   // a later block may read these endpoint bits: re-fit behind the walk (worker warp)
              else mark = 2;                   // nobody will: re-fit postponed to k_rdo_rehint
            }
            blocks[bi] = from_bits(best);
            modified[bi] = mark;
            uint64_t hi;
            ub::rdo_get_sel(best, first, total, key_lo, hi);   // (a re-fit does not touch the selector bits)
          }
        }
        ub::rdo_hist_set(hist, ch.hist_mask, first, key_lo, bi);
      }
      // the ring slot about to be overwritten belongs to block bi - 256: its re-fit, if queued, must have landed
      // (also keeps every block older than the ring final in global memory, where longer windows read it)
      for (;;) {
        const uint32_t tail = s_q_tail;
        if (tail == s_q_head || s_q_bi[tail % RDO_QUEUE] + 256u > bi) break;
        __nanosleep(100);
      }
      s_ring[bi & 255u] = from_bits(best);
      if (queued) {
        const uint32_t head = s_q_head;
        while (head - s_q_tail >= (uint32_t)RDO_QUEUE) __nanosleep(100);
        s_q_blk[head % RDO_QUEUE] = from_bits(best);
        s_q_bi[head % RDO_QUEUE] = bi;
        __threadfence_block();
        s_q_head = head + 1;
      }
    }
    if (bi + 1 < end) stash(b ^ 1);
    rdo_main_barrier();
  }
  if (tid == 0) { __threadfence_block(); s_q_done = 1; }
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

__global__ void k_set_u64(unsigned long long* p, unsigned long long v) { *p = v; }

// One mip level from the one above it: 2x2 box filter, per channel (a + b + c + d + 2) >> 2, source coordinates clamped to the
// level (so a 1-texel-wide or -high level averages the pair it has: (a + c + 1) >> 1).  SURVEY 8(d)'s harness definition; the
// CLI's float resampler (tools/ktx/command_create.cpp:1994-1996) is outside this path.  One thread per destination texel.
__global__ void __launch_bounds__(256) k_mip_down(const uint32_t* __restrict__ src, uint32_t sw, uint32_t sh, uint32_t* __restrict__ dst, uint32_t dw, uint32_t dh) {
  const uint64_t n = (uint64_t)dw * dh;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t y = (uint32_t)(i / dw), x = (uint32_t)(i - (uint64_t)y * dw);
    const uint32_t x0 = min(2 * x, sw - 1), x1 = min(2 * x + 1, sw - 1), y0 = min(2 * y, sh - 1), y1 = min(2 * y + 1, sh - 1);
    const uint32_t a = __ldg(src + (size_t)y0 * sw + x0), b = __ldg(src + (size_t)y0 * sw + x1);
    const uint32_t c = __ldg(src + (size_t)y1 * sw + x0), d = __ldg(src + (size_t)y1 * sw + x1);
    // two channels at a time in 16-bit lanes: sums <= 4 * 255 + 2 fit
    const uint32_t rb = ((a & 0x00FF00FFu) + (b & 0x00FF00FFu) + (c & 0x00FF00FFu) + (d & 0x00FF00FFu) + 0x00020002u) >> 2;
    const uint32_t ga = (((a >> 8) & 0x00FF00FFu) + ((b >> 8) & 0x00FF00FFu) + ((c >> 8) & 0x00FF00FFu) + ((d >> 8) & 0x00FF00FFu) + 0x00020002u) >> 2;
    dst[i] = (rb & 0x00FF00FFu) | ((ga & 0x00FF00FFu) << 8);
  }
}

#include "uastc_host.inc"

// ---------------------------------------------------------------------------------------------------------
// Allocation-free encode (no counterpart in the reference).
// b200_uastc_encode_image_dev launches k_encode_jobs with a stream-ordered scratch allocation per launch; callers that
// cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get the same kernel over a
// scratch that lives in the library's own device image: a file-scope array, one copy on every GPU that loads the library,
// held until the process exits.  Size: the tile counter + 1208 B x EJ_T blocks for each of at most B200_NOALLOC_MAX_CTAS
// resident tiles (one per SM of a B200).
constexpr unsigned B200_NOALLOC_MAX_CTAS = 148 * EJ_MIN_CTAS;   // every resident CTA of a B200
constexpr size_t B200_NOALLOC_SCRATCH_BYTES = EJ_SCRATCH_HEADER + (size_t)B200_NOALLOC_MAX_CTAS * EJ_SCRATCH_PER_CTA;   // ~131 MiB
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

namespace {
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[MAX_DEVICES] = {};   // device address of g_noalloc_scratch on each device (the symbol itself is a host-side stand-in)

// Device address of the scratch on the CURRENT device `dev`, looked up once per device with that device current.
int noalloc_scratch(int dev, void** out) {
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) return fail(B200_E_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch)", e);
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return B200_OK;
}
}  // namespace

extern "C" int b200_uastc_encode_image_dev_noalloc(const void* d_rgba, uint32_t width, uint32_t height, uint32_t row_pitch_bytes, void* d_out,
                                                   uint32_t pack_flags, void* stream, uint32_t* d_any_alpha) {
  if (!d_rgba || !d_out || !width || !height) return fail(B200_E_INVALID_VALUE, "bad image");
  const uint32_t pitch = row_pitch_bytes ? row_pitch_bytes : width * 4;
  if (pitch < width * 4) return fail(B200_E_INVALID_VALUE, "row pitch smaller than a row");
  if ((pitch & 3u) || ((uintptr_t)d_rgba & 3u)) return fail(B200_E_INVALID_VALUE, "row pitch / base must be 4-byte aligned");
  if ((uintptr_t)d_out & 15u) return fail(B200_E_INVALID_VALUE, "d_out must be 16-byte aligned");
  if ((pack_flags & 7u) >= 4u) return fail(B200_E_INVALID_VALUE, "b200_uastc_encode_image_dev_noalloc encodes levels 0-3 only (level 4 does not run the scratch-based kernel)");
  const uint32_t bx = (width + 3) / 4, by = (height + 3) / 4;
  const uint64_t n = (uint64_t)bx * by;
  if (n > B200_NOALLOC_MAX_BLOCKS) return fail(B200_E_INVALID_VALUE, "b200_uastc_encode_image_dev_noalloc: image exceeds B200_NOALLOC_MAX_BLOCKS (4194304 blocks = 8192 x 8192 texels)");
  const DeviceCtx* dc;
  int rc = ensure_device(cur_device(), &dc);
  if (rc) return rc;
  void* scratch = nullptr;
  rc = noalloc_scratch(dc->device, &scratch);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;   // NULL = the legacy default stream (no stream is created here)
  const uint64_t tiles = (n + EJ_T - 1) / EJ_T;
  const unsigned grid = (unsigned)std::min<uint64_t>(tiles, std::min<uint64_t>((uint64_t)dc->sm_count * EJ_MIN_CTAS, B200_NOALLOC_MAX_CTAS));
  CU(cudaMemsetAsync(scratch, 0, EJ_SCRATCH_HEADER, st));   // the tile counter must start at zero; the rest is written before it is read
  EncDescPack pack;
  pack.d[0] = {(const uint8_t*)d_rgba, (uint4*)d_out, width, height, pitch, bx, 0};
  k_encode_jobs<true, false><<<grid, EJ_T, sizeof(EjShared), st>>>(pack, nullptr, 1, n, pack_flags, d_any_alpha, (uint2*)scratch);
  g_launches++;
  CU(cudaGetLastError());
  if (!stream) CU(cudaStreamSynchronize(st));
  return B200_OK;
}

#if EJ_STAGE_CLOCKS
extern "C" int b200_uastc_debug_stage_clocks(unsigned long long* out20, int reset) {
  if (cudaMemcpyFromSymbol(out20, g_ej_clk, sizeof(unsigned long long) * 20) != cudaSuccess) return -1;
  if (reset) { unsigned long long z[20] = {0}; if (cudaMemcpyToSymbol(g_ej_clk, z, sizeof z) != cudaSuccess) return -1; }
  return 0;
}
#endif
