// uastc_unpack.cuh -- UASTC LDR 4x4 bitstream -> fields (restates unpack_uastc, reference
// external/basis_universal/transcoder/basisu_transcoder.cpp:14917-15374, in a register-friendly form).
//
// Layout choices that differ from the reference's unpacked struct:
//  * weights stay PACKED: one 128-bit string of `wbits`-wide entries in UASTC order (texel-major, plane-minor),
//    with the implicit zero MSB of every anchor weight re-inserted, so entry i sits at bit i*wbits;
//  * the partition pattern is a 2-bit-per-texel word, anchors are 4-bit fields of one word.
#pragma once
#include "ub_common.cuh"

namespace ub {

struct UnpackedUastc {
  uint32_t mode;
  uint32_t desc;       // ubt_uastc_mode_desc[mode]
  uint32_t pattern;    // common pattern index
  uint32_t ccs;        // dual-plane component selector
  uint32_t pat;        // ASTC subset of texel i = (pat >> 2i) & 3
  uint32_t anchors;    // ASTC anchor texel of subset s = (anchors >> 4s) & 15
  uint32_t solid;      // r | g<<8 | b<<16 | a<<24 (mode 8 only)
  uint32_t hints;      // raw hint bits (hint_bits wide), only when asked for
  uint8_t ep[18];      // BISE endpoint codes, RR GG BB (AA) per subset / LL AA
  Bits128 w;           // packed weights (see above)
};

UB_FN uint32_t spread_mask2(uint32_t m16) {  // 16-bit subset mask -> 2 bits per texel pattern (value 0/1)
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) r |= ((m16 >> i) & 1u) << (2 * i);
  return r;
}

// Returns false for an invalid block (reserved mode code, pattern index out of range): XC:14944, :15036-15049.
// KM >= 0: the caller already knows the mode (blocks binned by mode); every field offset then folds to a constant.
// TS: the per-CTA shared-memory tables are loaded (streaming transcode kernels): trit / quint bundles then decode by one look-up.
template <int KM, bool TS = false>
UB_FN bool unpack_uastc_t(const Bits128& blk, UnpackedUastc& u, bool want_hints) {
  const uint32_t mode = (KM >= 0) ? (uint32_t)KM : (uint32_t)UBT(uastc_huff_modes)[(uint32_t)blk.lo & 127u];
  if (KM < 0 && mode >= UASTC_TOTAL_MODES) return false;
  const uint32_t d = (KM >= 0) ? mode_desc_c(KM) : UBT(uastc_mode_desc)[mode];
  u.mode = mode;
  u.desc = d;
  u.pattern = 0;
  u.ccs = 0;
  u.pat = 0;
  u.anchors = 0;
  u.solid = 0;
  u.hints = 0;
  uint32_t ofs = md_huff_len(d);

  if (mode == UASTC_MODE_SOLID) {
    u.solid = bits_get(blk, ofs, 32);
    if (want_hints) u.hints = bits_get(blk, ofs + 32, 21);
    return true;
  }
  if (want_hints) u.hints = bits_get(blk, ofs, md_hint_bits(d));
  ofs += md_hint_bits(d);

  const uint32_t subsets = md_subsets(d);
  if (subsets == 3) {
    const uint32_t p = bits_get(blk, ofs, 4);
    ofs += 4;
    if (p >= 11) return false;
    u.pattern = p;
    u.pat = UBT(astc_pat3)[p];
    u.anchors = UBT(astc_anchor3)[p];
  } else if (subsets == 2) {
    const uint32_t p = bits_get(blk, ofs, 5);
    ofs += 5;
    u.pattern = p;
    if (mode == 7) {
      if (p >= 19) return false;
      u.pat = spread_mask2(UBT(astc_pat7)[p]);
      u.anchors = UBT(astc_anchor7)[p];
    } else {
      if (p >= 30) return false;
      u.pat = spread_mask2(UBT(astc_pat2)[p]);
      u.anchors = UBT(astc_anchor2)[p];
    }
  }

  const uint32_t planes = 1 + md_dual_plane(d);
  if (planes == 2) {
    if (mode == 17) u.ccs = 3;
    else { u.ccs = bits_get(blk, ofs, 2); ofs += 2; }
  }

  // ---- endpoints: trit/quint bundles first, then the plain bits of every value (XC:15085-15168)
  const uint32_t comps = md_comps(d);
  const uint32_t total_values = comps * 2 * subsets;
  const uint32_t er = md_endpoint_range(d);
  const uint32_t ep_bits = (KM >= 0) ? bise_c(er, 0) : UBT(astc_bise_range)[er * 3 + 0];
  const uint32_t has_trits = (KM >= 0) ? bise_c(er, 1) : UBT(astc_bise_range)[er * 3 + 1];
  const uint32_t has_quints = (KM >= 0) ? bise_c(er, 2) : UBT(astc_bise_range)[er * 3 + 2];

  uint32_t tq = 0;  // up to four bundles, 8 bits each
  uint32_t bundle = 0, mul = 0;
  if (has_trits | has_quints) {
    bundle = has_trits ? 5u : 3u;
    mul = has_trits ? 3u : 5u;
    const uint32_t total_tqs = (total_values + bundle - 1) / bundle;
    const uint32_t full_bits = has_trits ? 8u : 7u;
#pragma unroll
    for (uint32_t i = 0; i < 4; i++) {
      if (i < total_tqs) {
        uint32_t nb = full_bits;
        if (i == total_tqs - 1) {
          const uint32_t rem = total_values - (total_tqs - 1) * bundle;
          if (has_trits) { if (rem == 1) nb = 2; else if (rem == 2) nb = 4; else if (rem == 3) nb = 5; else if (rem == 4) nb = 7; }
          else { if (rem == 1) nb = 3; else if (rem == 2) nb = 5; }
        }
        tq |= bits_get(blk, ofs, nb) << (8 * i);
        ofs += nb;
      }
    }
  }
  {
    uint32_t accum = 0, left = 0, next = 0;
#pragma unroll
    for (uint32_t i = 0; i < 18; i++) {
      uint32_t v = 0;
      if (i < total_values) {
        v = bits_get(blk, ofs, ep_bits);
        ofs += ep_bits;
        if (bundle) {
#ifdef __CUDA_ARCH__
          if (TS) {   // accum = the bundle's digits, peeled by shifts
            if (!left) { const uint32_t b = (tq >> (8 * next)) & 255u; accum = has_trits ? ub_shared_tables.trits[b] : ub_shared_tables.quints[b & 127u]; next++; left = bundle; }
            const uint32_t dbits = has_trits ? 2u : 3u;
            v |= (accum & ((1u << dbits) - 1u)) << ep_bits;
            accum >>= dbits;
            left--;
          } else
#endif
          {
            if (!left) { accum = (tq >> (8 * next)) & 255u; next++; left = bundle; }
            const uint32_t q = accum / mul;
            v |= (accum - q * mul) << ep_bits;
            accum = q;
            left--;
          }
        }
      }
      u.ep[i] = (uint8_t)v;
    }
  }

  // ---- weights: everything from ofs to the end of the block; re-insert the anchors' implicit zero MSB
  const uint32_t wb = md_weight_bits(d);
  Bits128 w = bits_shr(blk, ofs);
  if (planes == 2) {
    w = bits_insert0(w, wb - 1);
    w = bits_insert0(w, 2 * wb - 1);
  } else {
    // anchors in ascending texel order; anchor texel 0 is always present
    uint32_t a0 = u.anchors & 15, a1 = (u.anchors >> 4) & 15, a2 = (u.anchors >> 8) & 15;
    if (subsets == 1) { w = bits_insert0(w, wb - 1); }
    else if (subsets == 2) {
      const uint32_t lo = a0 < a1 ? a0 : a1, hi = a0 < a1 ? a1 : a0;
      w = bits_insert0(w, lo * wb + wb - 1);
      w = bits_insert0(w, hi * wb + wb - 1);
    } else {
      // sort three
      uint32_t t;
      if (a0 > a1) { t = a0; a0 = a1; a1 = t; }
      if (a1 > a2) { t = a1; a1 = a2; a2 = t; }
      if (a0 > a1) { t = a0; a0 = a1; a1 = t; }
      w = bits_insert0(w, a0 * wb + wb - 1);
      w = bits_insert0(w, a1 * wb + wb - 1);
      w = bits_insert0(w, a2 * wb + wb - 1);
    }
  }
  // keep exactly 16*planes entries
  const uint32_t total_wbits = 16 * planes * wb;
  if (total_wbits <= 64) { w.lo &= lowmask64(total_wbits); w.hi = 0; }
  else w.hi &= lowmask64(total_wbits - 64);
  u.w = w;
  return true;
}

UB_FN bool unpack_uastc(const Bits128& blk, UnpackedUastc& u, bool want_hints) { return unpack_uastc_t<-1>(blk, u, want_hints); }

// weight entry i (UASTC order) of an unpacked block
UB_FN uint32_t uastc_weight(const UnpackedUastc& u, uint32_t i) {
  const uint32_t wb = md_weight_bits(u.desc);
  return bits_get(u.w, i * wb, wb);
}

// XOR mask that inverts the weights of every texel whose subset bit is set in `subset_sel` (bit s = invert subset s)
UB_FN Bits128 weight_invert_mask(const UnpackedUastc& u, uint32_t subset_sel) {
  const uint32_t wb = md_weight_bits(u.desc);
  const uint32_t per_texel = wb * (1 + md_dual_plane(u.desc));
  const uint64_t ones = lowmask64(per_texel);
  Bits128 m;
  m.lo = 0; m.hi = 0;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t s = (u.pat >> (2 * i)) & 3u;
    if ((subset_sel >> s) & 1u) bits_or(m, i * per_texel, ones);
  }
  return m;
}

}  // namespace ub
