// uastc_to_bc7.cuh -- UASTC -> BC7 block (restates transcode_uastc_to_bc7 + encode_bc7_block, reference
// external/basis_universal/transcoder/basisu_transcoder.cpp:15532-15647 (p-bit pickers), :15669-16161 (per-mode
// mapping) and :14292-14459 (BC7 bit packer)).  Output must be bit-exact with the reference.
#pragma once
#include "uastc_unpack.cuh"

namespace ub {

// The BC7 block under construction: selectors stay packed (ibits per texel) the whole way.
struct Bc7Build {
  uint32_t mode, partition, rotation;
  uint32_t lo[3], hi[3];   // r | g<<8 | b<<16 | a<<24 per subset
  uint32_t pbits;          // bit (2k) = low p-bit of subset k, bit (2k+1) = high p-bit
  uint64_t sel, asel;      // colour / alpha selectors, uniform width
};

UB_FN uint32_t px_c(uint32_t v, uint32_t c) { return (v >> (8 * c)) & 255u; }
UB_FN uint32_t px_set(uint32_t v, uint32_t c, uint32_t x) { return (v & ~(255u << (8 * c))) | (x << (8 * c)); }

// determine_unique_pbits (XC:15585-15647).  Each channel's quantised code and fp32 squared error depend only on the
// 8-bit value, the p-bit and the precision, so they come from tables (tools/gen_pbit_tables.py); what remains is the
// reference's fp32 summation order over channels and its strict-< choice between p = 0 and p = 1.
// u_lo/u_hi: 8-bit unquantised endpoints per channel; a channel the reference pins to 1.0f must be passed as 255.
// wide = true: 7+1 bits (BC7 modes 3 and 6); false: 5+1 bits (BC7 mode 7).
// One table entry per 8-bit channel value: fp32 error for p = 0 and p = 1, and both quantised codes.
// which: 0 = 7+1 bits (pb8_q), 1 = 5+1 bits (pb6_q), 2 = 6 bits + shared p-bit (pb7s_q)
template <bool TS = false>
UB_FN void pbit_entry(uint32_t which, uint32_t u, float& e0, float& e1, uint32_t& codes) {
  const uint32_t* tab = which == 0 ? UBT(pb8_q) : (which == 1 ? UBT(pb6_q) : UBT(pb7s_q));
#ifdef __CUDA_ARCH__
  const uint4 v = TS ? ub_shared_tables.pb[which][u] : __ldg((const uint4*)tab + u);
  e0 = __uint_as_float(v.x); e1 = __uint_as_float(v.y); codes = v.z;
#else
  e0 = bits_to_float(tab[u * 4]); e1 = bits_to_float(tab[u * 4 + 1]); codes = tab[u * 4 + 2];
#endif
}

template <bool TS = false>
UB_FN void unique_pbits(uint32_t total_comps, bool wide, uint32_t u_lo, uint32_t u_hi, uint32_t& out_lo, uint32_t& out_hi, uint32_t& pb) {
  const uint32_t tab = wide ? 0u : 1u;
  uint32_t ca[4], cb[4];   // per channel: code for p = 0 | code for p = 1 << 8
  float e_lo0 = 0, e_lo1 = 0, e_hi0 = 0, e_hi1 = 0;
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    float a0, a1, b0, b1;
    pbit_entry<TS>(tab, px_c(u_lo, c), a0, a1, ca[c]);
    pbit_entry<TS>(tab, px_c(u_hi, c), b0, b1, cb[c]);
    if (c < total_comps) { e_lo0 = f_add(e_lo0, a0); e_lo1 = f_add(e_lo1, a1); e_hi0 = f_add(e_hi0, b0); e_hi1 = f_add(e_hi1, b1); }
  }
  // gather the four channels' codes with byte permutes: x = {c0.p0, c1.p0, c0.p1, c1.p1}, y likewise for channels 2, 3
  const uint32_t xa = ub_prmt(ca[0], ca[1], 0x5140u), ya = ub_prmt(ca[2], ca[3], 0x5140u);
  const uint32_t xb = ub_prmt(cb[0], cb[1], 0x5140u), yb = ub_prmt(cb[2], cb[3], 0x5140u);
  const uint32_t q_lo0 = ub_prmt(xa, ya, 0x5410u), q_lo1 = ub_prmt(xa, ya, 0x7632u);
  const uint32_t q_hi0 = ub_prmt(xb, yb, 0x5410u), q_hi1 = ub_prmt(xb, yb, 0x7632u);
  // the reference takes p = 0 first (its error is always below the 1e+9 start value) and p = 1 only if strictly better
  const bool lo1 = e_lo1 < e_lo0, hi1 = e_hi1 < e_hi0;
  out_lo = lo1 ? q_lo1 : q_lo0;
  out_hi = hi1 ? q_hi1 : q_hi0;
  pb = (lo1 ? 1u : 0u) | (hi1 ? 2u : 0u);
}

// determine_shared_pbits (XC:15532-15582), BC7 mode 1 only (3 comps, 6+1 bits, alpha = 1.0f == 255)
template <bool TS = false>
UB_FN void shared_pbits(uint32_t u_lo, uint32_t u_hi, uint32_t& out_lo, uint32_t& out_hi, uint32_t& pb) {
  u_lo |= 0xFF000000u; u_hi |= 0xFF000000u;
  uint32_t ca[4], cb[4];
  float err0 = 0, err1 = 0;
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    float a0, a1, b0, b1;
    pbit_entry<TS>(2u, px_c(u_lo, c), a0, a1, ca[c]);
    pbit_entry<TS>(2u, px_c(u_hi, c), b0, b1, cb[c]);
    if (c < 3) { err0 = f_add(err0, f_add(a0, b0)); err1 = f_add(err1, f_add(a1, b1)); }
  }
  const uint32_t xa = ub_prmt(ca[0], ca[1], 0x5140u), ya = ub_prmt(ca[2], ca[3], 0x5140u);
  const uint32_t xb = ub_prmt(cb[0], cb[1], 0x5140u), yb = ub_prmt(cb[2], cb[3], 0x5140u);
  const uint32_t q_lo0 = ub_prmt(xa, ya, 0x5410u), q_lo1 = ub_prmt(xa, ya, 0x7632u);
  const uint32_t q_hi0 = ub_prmt(xb, yb, 0x5410u), q_hi1 = ub_prmt(xb, yb, 0x7632u);
  const bool p1 = err1 < err0;
  pb = p1 ? 3u : 0u;
  out_lo = p1 ? q_lo1 : q_lo0;
  out_hi = p1 ? q_hi1 : q_hi0;
}

// Re-pack the UASTC weights (entries of wb bits) of one plane into `ob`-bit BC7 selectors through a tiny value map.
// plane_stride = 1 (single plane) or 2; plane = 0/1.
//   kind: 0 identity, 1 = 5->4 bits, 2 = 2->4 bits, 3 = 3->4 bits, 4 = 1->2 bits (0 / 3)
// All 16 fields move at once (SIMD inside a 64-bit register) for every case but the 5-bit one: the per-texel loop was 11 % of the
// BC7 transcode kernel's instructions.
UB_FN uint64_t spread2to4(uint64_t x) {   // sixteen 2-bit fields -> the low 2 bits of sixteen nibbles
  x = (x | (x << 16)) & 0x0000FFFF0000FFFFull;
  x = (x | (x << 8)) & 0x00FF00FF00FF00FFull;
  x = (x | (x << 4)) & 0x0F0F0F0F0F0F0F0Full;
  x = (x | (x << 2)) & 0x3333333333333333ull;
  return x;
}
UB_FN uint64_t spread3to4(uint64_t x) {   // sixteen 3-bit fields (48 bits) -> the low 3 bits of sixteen nibbles
  x = (x & 0x0000000000FFFFFFull) | ((x & 0x0000FFFFFF000000ull) << 8);
  x = (x & 0x00000FFF00000FFFull) | ((x & 0x00FFF00000FFF000ull) << 4);
  x = (x & 0x003F003F003F003Full) | ((x & 0x0FC00FC00FC00FC0ull) << 2);
  x = (x & 0x0707070707070707ull) | ((x & 0x3838383838383838ull) << 1);
  return x;
}
UB_FN uint64_t even2(uint64_t x) {        // the even 2-bit fields of 32 -> sixteen packed 2-bit fields
  x &= 0x3333333333333333ull;
  x = (x | (x >> 2)) & 0x0F0F0F0F0F0F0F0Full;
  x = (x | (x >> 4)) & 0x00FF00FF00FF00FFull;
  x = (x | (x >> 8)) & 0x0000FFFF0000FFFFull;
  x = (x | (x >> 16)) & 0x00000000FFFFFFFFull;
  return x;
}
UB_FN uint64_t remap_selectors(const UnpackedUastc& u, uint32_t desc, uint32_t plane_stride, uint32_t plane, uint32_t ob, int kind) {
  const uint32_t wb = md_weight_bits(desc);
  if (plane_stride == 1 && kind == 0 && wb == ob) return u.w.lo;                       // same width: the packed weights are the selectors
  if (plane_stride == 1 && kind == 2 && wb == 2 && ob == 4) return spread2to4(u.w.lo) * 5ull;        // 0,5,10,15 per nibble, no carries
  if (plane_stride == 1 && kind == 3 && wb == 3 && ob == 4) {                          // 0,2,4,6,9,11,13,15 == 2w + (w >> 2)
    const uint64_t x = spread3to4(u.w.lo);
    return (x << 1) + ((x >> 2) & 0x1111111111111111ull);
  }
  if (plane_stride == 2 && kind == 0 && wb == 2 && ob == 2) return even2(u.w.lo >> (2 * plane));
  if (plane_stride == 2 && kind == 4 && wb == 1 && ob == 2) return ((u.w.lo >> plane) & 0x55555555ull) * 3ull;   // weight w at bit 2i -> w ? 3 : 0
  uint64_t r = 0;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) {
    uint32_t w = bits_get(u.w, (i * plane_stride + plane) * wb, wb);
    if (kind == 1) {
      // s_bc7_5_to_4 (XC:15756): 0,0,1,1,2,2,3,3,4,4,5,5,6,6,6,7,8,9,9,9,10,10,11,11,12,12,13,13,14,14,15,15
      const uint64_t t_lo = 0x7666554433221100ull, t_hi = 0xFFEEDDCCBBAA9998ull;
      w = (uint32_t)(((w < 16 ? t_lo : t_hi) >> ((w & 15u) * 4)) & 15u);
    } else if (kind == 2) {
      w = w * 5u;                                  // 0,5,10,15
    } else if (kind == 3) {
      w = (uint32_t)((0xFDB96420ull >> (w * 4)) & 15u);  // 0,2,4,6,9,11,13,15
    } else if (kind == 4) {
      w = w ? 3u : 0u;
    }
    r |= (uint64_t)w << (i * ob);
  }
  return r;
}

// expand a "texel belongs to subset k" test into an ob-bit-per-texel XOR mask
UB_FN uint64_t subset_mask64(uint32_t pat, uint32_t k, uint32_t ob) {
  const uint64_t ones = lowmask64(ob);
  uint64_t m = 0;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++)
    if (((pat >> (2 * i)) & 3u) == k) m |= ones << (i * ob);
  return m;
}

UB_FN uint32_t bc7_3_to_2(uint32_t p, uint32_t k) {  // bc7_convert_partition_index_3_to_2 (XC:13946)
  const uint32_t g = k >> 1;
  uint32_t r;
  if (g == 0) r = (p <= 1) ? 0u : 1u;
  else if (g == 1) r = (p == 0) ? 0u : 1u;
  else r = ((p == 0) || (p == 2)) ? 0u : 1u;
  return (k & 1u) ? 1u - r : r;
}

template <bool TS = false>
UB_FN uint32_t ep_unq(const UnpackedUastc& u, uint32_t er, uint32_t i) { return astc_unquant_s<TS>(er, u.ep[i]); }

// Gather the unquantised colour (RGB[A]) of one subset from the RR GG BB (AA) endpoint list.
template <bool TS = false>
UB_FN void subset_endpoints(const UnpackedUastc& u, uint32_t er, uint32_t base, uint32_t comps, uint32_t& lo, uint32_t& hi) {
  lo = 0; hi = 0;
#pragma unroll
  for (uint32_t c = 0; c < 4; c++)
    if (c < comps) {
      lo |= ep_unq<TS>(u, er, base + 2 * c) << (8 * c);
      hi |= ep_unq<TS>(u, er, base + 2 * c + 1) << (8 * c);
    }
}

// Per-mode mapping (XC:15669-16161).
template <int KM, bool TS = false>
UB_FN void uastc_to_bc7_build_t(const UnpackedUastc& u, Bc7Build& o) {
  const uint32_t mode = (KM >= 0) ? (uint32_t)KM : u.mode;
  const uint32_t desc = (KM >= 0) ? mode_desc_c(KM) : u.desc;
  const uint32_t er = md_endpoint_range(desc);
  const uint32_t comps = md_comps(desc);
  o.partition = 0; o.rotation = 0; o.pbits = 0; o.sel = 0; o.asel = 0;
#pragma unroll
  for (int k = 0; k < 3; k++) { o.lo[k] = 0; o.hi[k] = 0; }

  if (mode == UASTC_MODE_SOLID) {
    const uint32_t s = u.solid;
    uint32_t e0 = 0, e1 = 0;
#pragma unroll
    for (uint32_t c = 0; c < 4; c++) {
      e0 += UBT(bc7_mode6_opt)[px_c(s, c) * 2 + 0] & 0xFFFFu;
      e1 += UBT(bc7_mode6_opt)[px_c(s, c) * 2 + 1] & 0xFFFFu;
    }
    if (e0 > 0 && e1 > 0) {
      o.mode = 5;
#pragma unroll
      for (uint32_t c = 0; c < 3; c++) {
        const uint32_t t = UBT(bc7_mode5_opt)[px_c(s, c)];
        o.lo[0] |= ((t >> 16) & 255u) << (8 * c);
        o.hi[0] |= (t >> 24) << (8 * c);
      }
      o.lo[0] |= s & 0xFF000000u;
      o.hi[0] |= s & 0xFF000000u;
      o.sel = 0x55555555ull;  // index 1 in every 2-bit selector
    } else {
      o.mode = 6;
      const uint32_t bp = (e1 < e0) ? 1u : 0u;
#pragma unroll
      for (uint32_t c = 0; c < 4; c++) {
        const uint32_t t = UBT(bc7_mode6_opt)[px_c(s, c) * 2 + bp];
        o.lo[0] |= ((t >> 16) & 255u) << (8 * c);
        o.hi[0] |= (t >> 24) << (8 * c);
      }
      o.pbits = bp * 3u;
      o.sel = 0x5555555555555555ull;  // index 5 in every 4-bit selector
    }
    return;
  }

  switch (mode) {
    case 0: case 5: case 10: case 12: case 14: case 15: case 18: {
      o.mode = 6;
      uint32_t lo, hi;
      if (comps == 2) {
        const uint32_t l0 = ep_unq<TS>(u, er, 0), l1 = ep_unq<TS>(u, er, 1);
        lo = l0 * 0x010101u | (ep_unq<TS>(u, er, 2) << 24);
        hi = l1 * 0x010101u | (ep_unq<TS>(u, er, 3) << 24);
      } else {
        subset_endpoints<TS>(u, er, 0, comps, lo, hi);
        if (comps == 3) { lo |= 0xFF000000u; hi |= 0xFF000000u; }
      }
      uint32_t pb;
      unique_pbits<TS>(comps == 2 ? 4 : comps, true, lo, hi, o.lo[0], o.hi[0], pb);
      if (comps == 3) { o.lo[0] = px_set(o.lo[0], 3, 127); o.hi[0] = px_set(o.hi[0], 3, 127); }
      o.pbits = pb;
      const int kind = (mode == 18) ? 1 : (mode == 14) ? 2 : ((mode == 5) || (mode == 12)) ? 3 : 0;
      o.sel = remap_selectors(u, desc, 1, 0, 4, kind);
      break;
    }
    case 1: {
      o.mode = 3;
      uint32_t lo = 0, hi = 0;
#pragma unroll
      for (uint32_t c = 0; c < 3; c++) { lo |= (uint32_t)u.ep[2 * c] << (8 * c); hi |= (uint32_t)u.ep[2 * c + 1] << (8 * c); }
      uint32_t qlo, qhi, pb;
      unique_pbits<TS>(3, true, lo | 0xFF000000u, hi | 0xFF000000u, qlo, qhi, pb);
      o.lo[0] = o.lo[1] = qlo & 0xFFFFFFu;
      o.hi[0] = o.hi[1] = qhi & 0xFFFFFFu;
      o.pbits = pb | (pb << 2);
      o.sel = remap_selectors(u, desc, 1, 0, 2, 0);
      break;
    }
    case 2: {
      o.mode = 1;
      const uint32_t cpd = UBT(common2)[u.pattern];
      o.partition = cpd & 255u;
      const uint32_t inv = cpd >> 24;
#pragma unroll
      for (uint32_t s = 0; s < 2; s++) {
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (uint32_t c = 0; c < 3; c++) {
          lo |= ((uint32_t)u.ep[c * 2 + s * 6] * 17u) << (8 * c);
          hi |= ((uint32_t)u.ep[c * 2 + s * 6 + 1] * 17u) << (8 * c);
        }
        uint32_t qlo, qhi, pb;
        shared_pbits<TS>(lo, hi, qlo, qhi, pb);
        const uint32_t k = inv ? 1 - s : s;
        o.lo[k] = qlo & 0xFFFFFFu;
        o.hi[k] = qhi & 0xFFFFFFu;
        o.pbits |= (pb & 1u) << (2 * k);
      }
      o.sel = remap_selectors(u, desc, 1, 0, 3, 0);
      break;
    }
    case 3: {
      o.mode = 2;
      const uint32_t cpd = UBT(common3)[u.pattern];
      o.partition = cpd & 255u;
      const uint32_t perm = cpd >> 24;
#pragma unroll
      for (uint32_t s = 0; s < 3; s++) {
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (uint32_t c = 0; c < 3; c++) {
          lo |= ((ep_unq<TS>(u, er, c * 2 + s * 6) * 31u + 127u) / 255u) << (8 * c);
          hi |= ((ep_unq<TS>(u, er, c * 2 + 1 + s * 6) * 31u + 127u) / 255u) << (8 * c);
        }
        const uint32_t k = UBT(astc_to_bc7_perm)[perm * 3 + s];
        // k is data dependent: write through selects so lo[]/hi[] stay in registers
#pragma unroll
        for (uint32_t q = 0; q < 3; q++) if (q == k) { o.lo[q] = lo; o.hi[q] = hi; }
      }
      o.sel = remap_selectors(u, desc, 1, 0, 2, 0);
      break;
    }
    case 4: {
      o.mode = 3;
      const uint32_t cpd = UBT(common2)[u.pattern];
      o.partition = cpd & 255u;
      const uint32_t inv = cpd >> 24;
#pragma unroll
      for (uint32_t s = 0; s < 2; s++) {
        uint32_t lo, hi;
        subset_endpoints<TS>(u, er, s * 6, 3, lo, hi);
        uint32_t qlo, qhi, pb;
        unique_pbits<TS>(3, true, lo | 0xFF000000u, hi | 0xFF000000u, qlo, qhi, pb);
        const uint32_t k = inv ? 1 - s : s;
        o.lo[k] = px_set(qlo, 3, 127);
        o.hi[k] = px_set(qhi, 3, 127);
        o.pbits |= pb << (2 * k);
      }
      o.sel = remap_selectors(u, desc, 1, 0, 2, 0);
      break;
    }
    case 6: case 11: case 13: case 17: {
      o.mode = 5;
      o.rotation = (u.ccs + 1) & 3u;
      if (comps == 2) {
        const uint32_t l = (ep_unq<TS>(u, er, 0) * 127u + 127u) / 255u, h = (ep_unq<TS>(u, er, 1) * 127u + 127u) / 255u;
        o.lo[0] = l * 0x010101u | (ep_unq<TS>(u, er, 2) << 24);
        o.hi[0] = h * 0x010101u | (ep_unq<TS>(u, er, 3) << 24);
      } else {
#pragma unroll
        for (uint32_t ac = 0; ac < 4; ac++) {
          uint32_t bc = ac;
          if (ac == u.ccs) bc = 3; else if (ac == 3) bc = u.ccs;
          uint32_t l = 255, h = 255;
          if (ac < comps) { l = ep_unq<TS>(u, er, ac * 2); h = ep_unq<TS>(u, er, ac * 2 + 1); }
          if (bc < 3) { l = (l * 127u + 127u) / 255u; h = (h * 127u + 127u) / 255u; }
          o.lo[0] |= l << (8 * bc);
          o.hi[0] |= h << (8 * bc);
        }
      }
      const int kind = (mode == 13) ? 4 : 0;
      o.sel = remap_selectors(u, desc, 2, 0, 2, kind);
      o.asel = remap_selectors(u, desc, 2, 1, 2, kind);
      break;
    }
    case 7: {
      o.mode = 2;
      const uint32_t cpd = UBT(common7)[u.pattern];
      o.partition = cpd & 255u;
      const uint32_t kk = cpd >> 24;
#pragma unroll
      for (uint32_t bp = 0; bp < 3; bp++) {
        const uint32_t ap = bc7_3_to_2(bp, kk);
#pragma unroll
        for (uint32_t c = 0; c < 3; c++) {
          // ap is 0/1: select between the two subsets' endpoints without dynamic indexing
          const uint32_t l0 = u.ep[c * 2], l1 = u.ep[c * 2 + 6], h0 = u.ep[c * 2 + 1], h1 = u.ep[c * 2 + 7];
          o.lo[bp] |= ((astc_unquant_s<TS>(er, ap ? l1 : l0) * 31u + 127u) / 255u) << (8 * c);
          o.hi[bp] |= ((astc_unquant_s<TS>(er, ap ? h1 : h0) * 31u + 127u) / 255u) << (8 * c);
        }
      }
      o.sel = remap_selectors(u, desc, 1, 0, 2, 0);
      break;
    }
    default: {  // 9, 16
      o.mode = 7;
      const uint32_t cpd = UBT(common2)[u.pattern];
      o.partition = cpd & 255u;
      const uint32_t inv = cpd >> 24;
#pragma unroll
      for (uint32_t s = 0; s < 2; s++) {
        uint32_t lo, hi;
        if (comps == 2) {
          lo = ep_unq<TS>(u, er, s * 4) * 0x010101u | (ep_unq<TS>(u, er, 2 + s * 4) << 24);
          hi = ep_unq<TS>(u, er, 1 + s * 4) * 0x010101u | (ep_unq<TS>(u, er, 3 + s * 4) << 24);
        } else {
          subset_endpoints<TS>(u, er, s * 8, 4, lo, hi);
        }
        uint32_t qlo, qhi, pb;
        unique_pbits<TS>(4, false, lo, hi, qlo, qhi, pb);
        const uint32_t k = inv ? 1 - s : s;
        o.lo[k] = qlo; o.hi[k] = qhi;
        o.pbits |= pb << (2 * k);
      }
      o.sel = remap_selectors(u, desc, 1, 0, 2, 0);
      break;
    }
  }
}

UB_FN void uastc_to_bc7_build(const UnpackedUastc& u, Bc7Build& o) { uastc_to_bc7_build_t<-1>(u, o); }

// BC7 per-mode constants (XC:14262-14272), packed: [2:0]? -- kept as tiny switch-free tables in registers.
UB_FN uint32_t bc7_subsets(uint32_t m) { return (0x21112323u >> (4 * m)) & 15u; }      // 3,2,3,2,1,1,1,2
UB_FN uint32_t bc7_color_prec(uint32_t m) { return (0x57757564u >> (4 * m)) & 15u; }   // 4,6,5,7,5,7,7,5
UB_FN uint32_t bc7_alpha_prec(uint32_t m) { return (0x57860000u >> (4 * m)) & 15u; }   // 0,0,0,0,6,8,7,5
UB_FN uint32_t bc7_index_bits(uint32_t m) { return (0x24222233u >> (4 * m)) & 15u; }   // 3,3,2,2,2,2,4,2

// encode_bc7_block (XC:14292-14459) for the BC7 modes UASTC maps to (1,2,3,5,6,7; index_selector is always 0).
// BM >= 0: the BC7 mode is a compile-time constant (the mode-specialised transcoders): every field position folds, and the block
// is assembled in 32-bit words (BitWriter32).  BM < 0: the mode is o.mode.
template <int BM, typename WRITER>
UB_FN Bits128 bc7_pack_t(Bc7Build& o) {
  const uint32_t m = (BM >= 0) ? (uint32_t)BM : o.mode;
  const uint32_t subsets = bc7_subsets(m);
  const uint32_t ib = bc7_index_bits(m);
  uint32_t pat = 0;
  uint32_t anchor[3] = {0, 0, 0};
  if (subsets == 2) { pat = spread_mask2(UBT(bc7_part2)[o.partition]); anchor[1] = UBT(bc7_anchor2)[o.partition]; }
  else if (subsets == 3) { pat = UBT(bc7_part3)[o.partition]; anchor[1] = UBT(bc7_anchor3a)[o.partition]; anchor[2] = UBT(bc7_anchor3b)[o.partition]; }

  const bool sep_alpha = (m == 5);
  const bool shared_p = (m == 1);
#pragma unroll
  for (uint32_t k = 0; k < 3; k++) {
    if (k < subsets) {
      const uint32_t a = anchor[k];
      if ((o.sel >> (a * ib + ib - 1)) & 1ull) {
        o.sel ^= subset_mask64(pat, k, ib);
        if (sep_alpha) {
          const uint32_t l = o.lo[k], h = o.hi[k];
          o.lo[k] = (l & 0xFF000000u) | (h & 0xFFFFFFu);
          o.hi[k] = (h & 0xFF000000u) | (l & 0xFFFFFFu);
        } else {
          const uint32_t t = o.lo[k]; o.lo[k] = o.hi[k]; o.hi[k] = t;
        }
        if (!shared_p) {
          const uint32_t pb = (o.pbits >> (2 * k)) & 3u;
          o.pbits = (o.pbits & ~(3u << (2 * k))) | ((((pb & 1u) << 1) | (pb >> 1)) << (2 * k));
        }
      }
      if (sep_alpha) {
        if ((o.asel >> (a * 2 + 1)) & 1ull) {
          o.asel ^= subset_mask64(pat, k, 2);
          const uint32_t l = o.lo[k], h = o.hi[k];
          o.lo[k] = (h & 0xFF000000u) | (l & 0xFFFFFFu);
          o.hi[k] = (l & 0xFF000000u) | (h & 0xFFFFFFu);
        }
      }
    }
  }

  WRITER bw;
  bw.init();
  bw.put(1u << m, m + 1);
  if (m == 5) bw.put(o.rotation, 2);
  if (subsets > 1) bw.put(o.partition, (m == 0) ? 4 : 6);
  const uint32_t comps = (m >= 4) ? 4u : 3u;
  const uint32_t cp = bc7_color_prec(m), ap = bc7_alpha_prec(m);
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    if (c < comps) {
      const uint32_t nb = (c == 3) ? ap : cp;
#pragma unroll
      for (uint32_t k = 0; k < 3; k++)
        if (k < subsets) { bw.put(px_c(o.lo[k], c), nb); bw.put(px_c(o.hi[k], c), nb); }
    }
  }
  if (m == 1 || m == 3 || m == 6 || m == 7) {
#pragma unroll
    for (uint32_t k = 0; k < 3; k++)
      if (k < subsets) {
        bw.put((o.pbits >> (2 * k)) & 1u, 1);
        if (!shared_p) bw.put((o.pbits >> (2 * k + 1)) & 1u, 1);
      }
  }
  // selectors: drop each anchor's (now zero) MSB, highest position first so lower positions stay valid
  {
    uint32_t a1 = anchor[1], a2 = anchor[2];
    uint64_t s = o.sel;
    uint32_t total = 16 * ib;
    if (subsets == 3) {
      const uint32_t hi = a1 > a2 ? a1 : a2, lo = a1 > a2 ? a2 : a1;
      s = u64_delete_bit(s, hi * ib + ib - 1);
      s = u64_delete_bit(s, lo * ib + ib - 1);
      total -= 2;
    } else if (subsets == 2) {
      s = u64_delete_bit(s, a1 * ib + ib - 1);
      total -= 1;
    }
    s = u64_delete_bit(s, ib - 1);
    total -= 1;
    bw.put64(s, total);
  }
  if (sep_alpha) {
    const uint64_t s = u64_delete_bit(o.asel, 1);
    bw.put64(s, 31);
  }
  return bw.bits();
}
UB_FN Bits128 bc7_pack(Bc7Build& o) { return bc7_pack_t<-1, BitWriter>(o); }

// BC7 mode a UASTC mode transcodes to (XC:15669-16161); -1 = solid colour blocks: mode 5 or 6, decided per block
UB_CONSTEXPR int bc7_mode_of_uastc(int m) {
  constexpr int t[19] = {6, 3, 1, 2, 3, 6, 5, 2, -1, 7, 6, 5, 6, 5, 6, 6, 7, 5, 6};
  return (m < 0 || m > 18) ? -1 : t[m];
}

// Full path for one block.  Returns false on an invalid UASTC block (output then zeroed).
template <int KM, bool TS = false>
UB_FN bool transcode_block_to_bc7_t(const Bits128& src, Bits128& dst) {
  UnpackedUastc u;
  if (!unpack_uastc_t<KM, TS>(src, u, false)) { dst.lo = 0; dst.hi = 0; return false; }
  Bc7Build o;
  uastc_to_bc7_build_t<KM, TS>(u, o);
  constexpr int BM = bc7_mode_of_uastc(KM);
  if (KM < 0) dst = bc7_pack(o);
  else if (BM >= 0) dst = bc7_pack_t<BM, BitWriter32>(o);
  else dst = (o.mode == 5) ? bc7_pack_t<5, BitWriter32>(o) : bc7_pack_t<6, BitWriter32>(o);
  return true;
}
UB_FN bool transcode_block_to_bc7(const Bits128& src, Bits128& dst) { return transcode_block_to_bc7_t<-1>(src, dst); }

}  // namespace ub
