// ub_common.cuh -- shared definitions for the B200 UASTC module.
//
// Every algorithmic routine in csrc/*.cuh is written once as a UB_FN function.  nvcc compiles it as
// __device__ code for sm_100a (the product).  g++ can compile the very same headers as plain inline C++
// (oracle/port_host.cpp), which is how the restated algorithm is diffed against the unmodified reference
// on a box without a GPU.  The host build is test infrastructure only; the product library contains no
// host implementation of the hot path.
#pragma once
#include <stdint.h>
#include <string.h>
#include <math.h>

#ifdef __CUDACC__
#define UB_FN __device__ __forceinline__
#define UB_MFN __device__ __forceinline__
#define UB_FN_NOINLINE __device__ __noinline__
#define UB_BIG __device__ __noinline__   /* large encoder stages: one out-of-line copy, called from every mode */
#define UB_TABLE(type, name, n) static __device__ const type ubt_##name[n]
#define UB_TABLE_ALIGNED16(type, name, n) static __device__ __align__(16) const type ubt_##name[n]
#define UB_CONSTEXPR __host__ __device__ constexpr
#else
#define UB_FN static inline
#define UB_MFN inline
#define UB_FN_NOINLINE static
#define UB_BIG static
#define UB_TABLE(type, name, n) static const type ubt_##name[n]
#define UB_TABLE_ALIGNED16(type, name, n) alignas(16) static const type ubt_##name[n]
#define UB_CONSTEXPR constexpr
#endif
#define UBT(name) ubt_##name
// Stage boundary inside the encoder: when a CTA runs the encoder in lock-step (EncCfg::cta_sync, uniform over the CTA)
// its warps wait for each other here, so they stay within the same stage's code.  Must be reached by every thread.
#ifdef __CUDA_ARCH__
#define UB_STAGE_SYNC(cfg) do { if ((cfg).cta_sync) __syncthreads(); } while (0)
#else
#define UB_STAGE_SYNC(cfg) do { } while (0)
#endif

namespace ub {

// ---- IEEE single ops with no contraction.  The reference is built with -ffp-contract=off (cmake/fp-settings.cmake:26)
// so every float op rounds once; on the device we spell the rounding mode so ptxas can never fuse.
#ifdef __CUDA_ARCH__
UB_FN float f_add(float a, float b) { return __fadd_rn(a, b); }
UB_FN float f_sub(float a, float b) { return __fsub_rn(a, b); }
UB_FN float f_mul(float a, float b) { return __fmul_rn(a, b); }
UB_FN float f_div(float a, float b) { return __fdiv_rn(a, b); }
UB_FN float f_sqrt(float a) { return __fsqrt_rn(a); }
UB_FN double d_add(double a, double b) { return __dadd_rn(a, b); }
UB_FN double d_sub(double a, double b) { return __dsub_rn(a, b); }
UB_FN double d_mul(double a, double b) { return __dmul_rn(a, b); }
UB_FN double d_div(double a, double b) { return __ddiv_rn(a, b); }
#else
UB_FN float f_add(float a, float b) { return a + b; }
UB_FN float f_sub(float a, float b) { return a - b; }
UB_FN float f_mul(float a, float b) { return a * b; }
UB_FN float f_div(float a, float b) { return a / b; }
UB_FN float f_sqrt(float a) { return sqrtf(a); }
UB_FN double d_add(double a, double b) { return a + b; }
UB_FN double d_sub(double a, double b) { return a - b; }
UB_FN double d_mul(double a, double b) { return a * b; }
UB_FN double d_div(double a, double b) { return a / b; }
#endif

UB_FN float bits_to_float(uint32_t b) {
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f; memcpy(&f, &b, 4); return f;
#endif
}

UB_FN int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
UB_FN uint32_t minu(uint32_t a, uint32_t b) { return a < b ? a : b; }
UB_FN uint32_t maxu(uint32_t a, uint32_t b) { return a > b ? a : b; }

// ---- a 128-bit little-endian bit string (one compressed block)
struct Bits128 {
  uint64_t lo, hi;
};

UB_FN uint64_t lowmask64(uint32_t n) { return n >= 64 ? ~0ull : ((1ull << n) - 1ull); }

// bits [ofs, ofs+n) of b, n <= 32, zero-filled beyond bit 127
UB_FN uint32_t bits_get(const Bits128& b, uint32_t ofs, uint32_t n) {
  uint64_t v;
  if (ofs >= 128) v = 0;
  else if (ofs >= 64) v = b.hi >> (ofs - 64);
  else if (ofs == 0) v = b.lo;
  else v = (b.lo >> ofs) | (b.hi << (64 - ofs));
  return (uint32_t)v & (uint32_t)lowmask64(n);
}

// 64 bits starting at ofs, zero-filled beyond bit 127
UB_FN uint64_t bits_get64(const Bits128& b, uint32_t ofs) {
  if (ofs >= 128) return 0;
  if (ofs >= 64) return b.hi >> (ofs - 64);
  if (ofs == 0) return b.lo;
  return (b.lo >> ofs) | (b.hi << (64 - ofs));
}

UB_FN Bits128 bits_shr(const Bits128& b, uint32_t s) {
  Bits128 r;
  if (s == 0) return b;
  if (s >= 128) { r.lo = 0; r.hi = 0; }
  else if (s >= 64) { r.lo = b.hi >> (s - 64); r.hi = 0; }
  else { r.lo = (b.lo >> s) | (b.hi << (64 - s)); r.hi = b.hi >> s; }
  return r;
}

// OR `val` (n <= 64 significant bits) in at bit ofs; bits that fall past 127 are dropped
UB_FN void bits_or(Bits128& b, uint32_t ofs, uint64_t val) {
  if (ofs >= 128) return;
  if (ofs >= 64) { b.hi |= val << (ofs - 64); return; }
  b.lo |= val << ofs;
  if (ofs) b.hi |= val >> (64 - ofs);
}

struct BitWriter {
  Bits128 b;
  uint32_t pos;
  UB_MFN void init() { b.lo = 0; b.hi = 0; pos = 0; }
  UB_MFN void put(uint64_t val, uint32_t n) { bits_or(b, pos, val); pos += n; }
  UB_MFN void put64(uint64_t val, uint32_t n) { put(val, n); }
  UB_MFN Bits128 bits() const { return b; }
};

// One compressed block assembled in four 32-bit words.  For code whose field positions are compile-time constants after
// unrolling (the mode-specialised packers): every put then is one shift + one OR on a 32-bit register (two when the field
// straddles a word), where the 64-bit BitWriter spends a shift pair and an OR pair per field.
struct BitWriter32 {
  uint32_t w0, w1, w2, w3, pos;
  UB_MFN void init() { w0 = 0; w1 = 0; w2 = 0; w3 = 0; pos = 0; }
  UB_MFN void or_word(uint32_t k, uint32_t v) { if (k == 0) w0 |= v; else if (k == 1) w1 |= v; else if (k == 2) w2 |= v; else if (k == 3) w3 |= v; }
  UB_MFN void put(uint32_t val, uint32_t n) {   // n <= 32, val < 2^n; bits past 127 are dropped
    const uint32_t k = pos >> 5, sh = pos & 31u;
    or_word(k, val << sh);
    if (sh && sh + n > 32u) or_word(k + 1, val >> (32u - sh));
    pos += n;
  }
  UB_MFN void put64(uint64_t val, uint32_t n) {  // n <= 64
    put((uint32_t)val, n < 32u ? n : 32u);
    if (n > 32u) put((uint32_t)(val >> 32), n - 32u);
  }
  UB_MFN Bits128 bits() const { Bits128 b; b.lo = (uint64_t)w0 | ((uint64_t)w1 << 32); b.hi = (uint64_t)w2 | ((uint64_t)w3 << 32); return b; }
};

// __byte_perm: result byte i = byte (sel >> 4i) & 7 of the eight bytes {a0..a3, b0..b3}
UB_FN uint32_t ub_prmt(uint32_t a, uint32_t b, uint32_t sel) {
#ifdef __CUDA_ARCH__
  return __byte_perm(a, b, sel);
#else
  const uint64_t v = (uint64_t)a | ((uint64_t)b << 32);
  uint32_t r = 0;
  for (int i = 0; i < 4; i++) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7u))) & 255u) << (8 * i);
  return r;
#endif
}

// insert a zero bit at position p (bits >= p move up by one)
UB_FN Bits128 bits_insert0(const Bits128& b, uint32_t p) {
  Bits128 r;
  if (p >= 64) {
    const uint64_t m = lowmask64(p - 64);
    r.lo = b.lo;
    r.hi = (b.hi & m) | ((b.hi & ~m) << 1);
  } else {
    const uint64_t m = lowmask64(p);
    r.lo = (b.lo & m) | ((b.lo & ~m) << 1);
    r.hi = (b.hi << 1) | (b.lo >> 63);
  }
  return r;
}

// delete bit p of a 64-bit word (bits above move down)
UB_FN uint64_t u64_delete_bit(uint64_t x, uint32_t p) {
  const uint64_t m = lowmask64(p);
  return (x & m) | ((x >> 1) & ~m);
}

UB_FN uint64_t brev64(uint64_t x) {
#ifdef __CUDA_ARCH__
  return __brevll(x);
#else
  x = ((x >> 1) & 0x5555555555555555ull) | ((x & 0x5555555555555555ull) << 1);
  x = ((x >> 2) & 0x3333333333333333ull) | ((x & 0x3333333333333333ull) << 2);
  x = ((x >> 4) & 0x0F0F0F0F0F0F0F0Full) | ((x & 0x0F0F0F0F0F0F0F0Full) << 4);
  x = ((x >> 8) & 0x00FF00FF00FF00FFull) | ((x & 0x00FF00FF00FF00FFull) << 8);
  x = ((x >> 16) & 0x0000FFFF0000FFFFull) | ((x & 0x0000FFFF0000FFFFull) << 16);
  return (x >> 32) | (x << 32);
#endif
}

#include "uastc_tables.inc"
#include "uastc_pbit_tables.inc"

// ---- decoded fields of ubt_uastc_mode_desc (see tools/gen_tables.py for the packing)
UB_FN uint32_t md_weight_bits(uint32_t d) { return d & 7; }
UB_FN uint32_t md_endpoint_range(uint32_t d) { return (d >> 3) & 31; }
UB_FN uint32_t md_subsets(uint32_t d) { return (d >> 8) & 3; }
UB_FN uint32_t md_dual_plane(uint32_t d) { return (d >> 10) & 1; }
UB_FN uint32_t md_comps(uint32_t d) { return (d >> 11) & 7; }
UB_FN uint32_t md_hint_bits(uint32_t d) { return (d >> 14) & 31; }
UB_FN uint32_t md_huff_len(uint32_t d) { return (d >> 19) & 15; }
UB_FN uint32_t md_huff_code(uint32_t d) { return (d >> 23) & 127; }


// ---- compile-time copies of the two tiny tables the mode-specialised transcoders fold on (checked against the
// generated tables by tests/test_oracle_cpu.py::test_constexpr_tables_match)
UB_CONSTEXPR uint32_t mode_desc_c(int m) {
  constexpr uint32_t t[19] = {0x00a3d99cu,0x1ab3d9a2u,0x0eabda43u,0x01abdb3au,0x09abda62u,0x05abd9a3u,0x0dabdd92u,0x03abda62u,0x0ba82000u,0x07ade242u,0x011c616cu,0x0014656au,0x031c619bu,0x0fade5a1u,0x06ade1a2u,0x02bdd1a4u,0x0ab5d2a2u,0x12b5d5a2u,0x04a3d95du};
  return t[m < 0 ? 0 : m];
}
UB_CONSTEXPR uint32_t bise_c(uint32_t range, uint32_t k) {  // k: 0 = bits, 1 = trits, 2 = quints
  constexpr uint8_t t[63] = {1,0,0,0,1,0,2,0,0,0,0,1,1,1,0,3,0,0,1,0,1,2,1,0,4,0,0,2,0,1,3,1,0,5,0,0,3,0,1,4,1,0,6,0,0,4,0,1,5,1,0,7,0,0,5,0,1,6,1,0,8,0,0};
  return t[range * 3 + k];
}

// Table space.  TS = true: the hot look-up tables of the streaming transcode kernels (ASTC endpoint unquantisation,
// BC7 p-bit entries) are read from a per-CTA copy in shared memory (shared_tables_load) -- random 2- and 16-byte
// look-ups cost a few shared-memory wavefronts there instead of one L1 tag look-up per distinct line.  TS = false
// (the default, and the only choice on the host): the global tables.
#ifdef __CUDACC__
struct SharedTables {
  uint8_t unq[21 * 256];    // low byte of ubt_astc_unquant: the unquantised value
  uint4 pb[3][256];         // ubt_pb8_q, ubt_pb6_q, ubt_pb7s_q
  uint16_t trits[256];      // BISE bundle -> its five base-3 digits, 2 bits each (digit k at bits 2k)
  uint16_t quints[128];     // BISE bundle -> its three base-5 digits, 3 bits each
};
__shared__ SharedTables ub_shared_tables;   // only kernels that reference it get the 23 KB allocation
__device__ __forceinline__ void shared_tables_load(uint32_t tid, uint32_t nthreads) {   // followed by a CTA barrier at the caller
  for (uint32_t i = tid; i < 21 * 256 / 4; i += nthreads) {
    const uint2 v = ((const uint2*)UBT(astc_unquant))[i];   // four 16-bit entries -> their four low bytes
    ((uint32_t*)ub_shared_tables.unq)[i] = (v.x & 0xFFu) | ((v.x >> 8) & 0xFF00u) | ((v.y & 0xFFu) << 16) | ((v.y & 0xFF0000u) << 8);
  }
  for (uint32_t i = tid; i < 256; i += nthreads) {
    uint32_t v = i, t = 0, q = 0;
    for (uint32_t k = 0; k < 5; k++) { t |= (v % 3u) << (2 * k); v /= 3u; }   // the reference peels digits with % and / (XC:15142-15168)
    v = i;
    for (uint32_t k = 0; k < 3; k++) { q |= (v % 5u) << (3 * k); v /= 5u; }
    ub_shared_tables.trits[i] = (uint16_t)t;
    if (i < 128) ub_shared_tables.quints[i] = (uint16_t)q;
    ub_shared_tables.pb[0][i] = ((const uint4*)UBT(pb8_q))[i];
    ub_shared_tables.pb[1][i] = ((const uint4*)UBT(pb6_q))[i];
    ub_shared_tables.pb[2][i] = ((const uint4*)UBT(pb7s_q))[i];
  }
}
#endif

// TS = true is used by the mode-specialised transcoders only, where `range` is a compile-time constant: the ranges without
// trits or quints unquantise by bit replication (range 20: 8 bits, identity; 8: 4 bits; 11: 5 bits) and need no look-up
// (checked against the table by tests/test_oracle_cpu.py::test_bit_only_unquant_identities).
template <bool TS = false>
UB_FN uint32_t astc_unquant_s(uint32_t range, uint32_t code) {
#if defined(__CUDA_ARCH__)
  if (TS) {
    if (range == 20) return code;
    if (range == 8) return code * 17u;
    if (range == 11) return (code << 3) | (code >> 2);
    return ub_shared_tables.unq[range * 256 + code];
  }
#endif
  return UBT(astc_unquant)[range * 256 + code] & 0xFF;
}
UB_FN uint32_t astc_unquant(uint32_t range, uint32_t code) { return astc_unquant_s<false>(range, code); }

const uint32_t UASTC_MODE_SOLID = 8;
const uint32_t UASTC_TOTAL_MODES = 19;

}  // namespace ub
