// uastc_decode.cuh -- block -> 16 RGBA texels, for the encoder's candidate scoring and for PSNR reporting.
//  * decode_uastc_pixels restates unpack_uastc(struct -> pixels), reference
//    external/basis_universal/transcoder/basisu_transcoder.cpp:15378-15514 (+ astc_interpolate, basisu_transcoder_uastc.h:76).
//  * decode_bc7_pixels restates bc7u::unpack_bc7 for the BC7 modes UASTC transcodes to (1,2,3,5,6,7), XC:29298-29640.
#pragma once
#include "uastc_to_bc7.cuh"

namespace ub {

// One encoder candidate before bit packing (the reference's uastc_encode_results / astc_block_desc).
struct Candidate {
  uint32_t mode, pattern, ccs;
  uint8_t ep[18];  // BISE codes, RR GG BB (AA) per ASTC subset, or LL AA
  uint8_t w[32];   // weights, texel-major / plane-minor, full width (anchor MSBs not yet normalised)
};

UB_FN uint32_t astc_interp(uint32_t l, uint32_t h, uint32_t w) {  // astc_interpolate, srgb = false
  l = (l << 8) | l; h = (h << 8) | h;
  return ((l * (64 - w) + h * w + 32) >> 6) >> 8;
}

UB_FN uint32_t sel_weight_bits(uint32_t wb, uint32_t l) {  // g_astc_weight_tables[wb][l] (XC:15376)
  switch (wb) {
    case 1: return l ? 64u : 0u;
    case 2: return UBT(bc7_weights2)[l];
    case 3: return UBT(bc7_weights3)[l];
    case 4: return UBT(astc_weights4)[l];
    default: return UBT(astc_weights5)[l];
  }
}

UB_FN uint32_t cand_pat(const Candidate& c, uint32_t d) {
  const uint32_t subsets = md_subsets(d);
  if (subsets == 3) return UBT(astc_pat3)[c.pattern];
  if (subsets == 2) return spread_mask2(c.mode == 7 ? UBT(astc_pat7)[c.pattern] : UBT(astc_pat2)[c.pattern]);
  return 0;
}

UB_FN void decode_uastc_pixels(const Candidate& c, uint32_t* out) {
  const uint32_t d = UBT(uastc_mode_desc)[c.mode];
  const uint32_t subsets = md_subsets(d), comps = md_comps(d), er = md_endpoint_range(d), wb = md_weight_bits(d);
  const uint32_t planes = 1 + md_dual_plane(d);
  const uint32_t levels = 1u << wb;
  uint32_t e0[3], e1[3];
  for (uint32_t s = 0; s < subsets; s++) {
    if (comps == 2) {
      const uint32_t ll = astc_unquant(er, c.ep[s * 4 + 0]), lh = astc_unquant(er, c.ep[s * 4 + 1]);
      const uint32_t al = astc_unquant(er, c.ep[s * 4 + 2]), ah = astc_unquant(er, c.ep[s * 4 + 3]);
      e0[s] = ll * 0x010101u | (al << 24);
      e1[s] = lh * 0x010101u | (ah << 24);
    } else {
      uint32_t a = 0, b = 0;
      for (uint32_t k = 0; k < 4; k++) {
        uint32_t l = 255, h = 255;
        if (k < comps) { l = astc_unquant(er, c.ep[s * comps * 2 + k * 2]); h = astc_unquant(er, c.ep[s * comps * 2 + k * 2 + 1]); }
        a |= l << (8 * k); b |= h << (8 * k);
      }
      e0[s] = a; e1[s] = b;
    }
  }
  uint32_t colors[3][32];
#pragma unroll 1
  for (uint32_t s = 0; s < subsets; s++)
#pragma unroll 1
    for (uint32_t l = 0; l < levels; l++) {
      const uint32_t w = sel_weight_bits(wb, l);
      uint32_t v = 0;
      for (uint32_t k = 0; k < 4; k++) {
        uint32_t x;
        if (comps == 2 || k < comps) x = astc_interp(px_c(e0[s], k), px_c(e1[s], k), w); else x = 255;
        v |= x << (8 * k);
      }
      colors[s][l] = v;
    }
  const uint32_t pat = cand_pat(c, d);
#pragma unroll 1
  for (uint32_t i = 0; i < 16; i++) {
    if (planes == 1) {
      out[i] = colors[(pat >> (2 * i)) & 3u][c.w[i]];
    } else {
      const uint32_t c0 = colors[0][c.w[i * 2]], c1 = colors[0][c.w[i * 2 + 1]];
      const uint32_t m = 255u << (8 * c.ccs);
      out[i] = (c0 & ~m) | (c1 & m);
    }
  }
}

// BC7 block -> texels, modes 1,2,3,5,6,7 (anything else returns false; the UASTC transcoder never emits them)
UB_FN uint32_t bc7_dq(uint32_t v, uint32_t bits) { v <<= (8 - bits); return v | (v >> bits); }
UB_FN uint32_t bc7_w(uint32_t bits, uint32_t i) {
  return bits == 2 ? UBT(bc7_weights2)[i] : (bits == 3 ? UBT(bc7_weights3)[i] : UBT(bc7_weights4)[i]);
}
UB_FN uint32_t bc7_lerp(uint32_t l, uint32_t h, uint32_t w) { return (l * (64 - w) + h * w + 32) >> 6; }

UB_FN bool decode_bc7_pixels(const Bits128& b, uint32_t* out) {
  uint32_t mode = 0;
  while (mode < 8 && !((b.lo >> mode) & 1ull)) mode++;
  if (mode == 0 || mode == 4 || mode >= 8) return false;
  uint32_t pos = mode + 1;
  if (mode == 5) {
    const uint32_t rot = bits_get(b, pos, 2); pos += 2;
    uint32_t e[2][4];
    for (uint32_t c = 0; c < 4; c++)
      for (uint32_t k = 0; k < 2; k++) { const uint32_t nb = (c == 3) ? 8u : 7u; e[k][c] = bc7_dq(bits_get(b, pos, nb), nb); pos += nb; }
    uint32_t cw[16], aw[16];
    for (uint32_t i = 0; i < 16; i++) { const uint32_t nb = i ? 2u : 1u; cw[i] = bits_get(b, pos, nb); pos += nb; }
    for (uint32_t i = 0; i < 16; i++) { const uint32_t nb = i ? 2u : 1u; aw[i] = bits_get(b, pos, nb); pos += nb; }
    for (uint32_t i = 0; i < 16; i++) {
      uint32_t p[4];
      for (uint32_t c = 0; c < 3; c++) p[c] = bc7_lerp(e[0][c], e[1][c], UBT(bc7_weights2)[cw[i]]);
      p[3] = bc7_lerp(e[0][3], e[1][3], UBT(bc7_weights2)[aw[i]]);
      if (rot >= 1) { const uint32_t t = p[3]; p[3] = p[rot - 1]; p[rot - 1] = t; }
      out[i] = p[0] | (p[1] << 8) | (p[2] << 16) | (p[3] << 24);
    }
    return true;
  }
  const uint32_t subsets = bc7_subsets(mode), ib = bc7_index_bits(mode);
  uint32_t part = 0;
  if (subsets > 1) { part = bits_get(b, pos, 6); pos += 6; }
  const uint32_t comps = (mode >= 4) ? 4u : 3u;
  const uint32_t cp = bc7_color_prec(mode), ap = bc7_alpha_prec(mode);
  uint32_t e[6][4];
  for (uint32_t c = 0; c < comps; c++)
    for (uint32_t k = 0; k < subsets * 2; k++) { const uint32_t nb = (c == 3) ? ap : cp; e[k][c] = bits_get(b, pos, nb); pos += nb; }
  const bool has_p = (mode == 1 || mode == 3 || mode == 6 || mode == 7);
  uint32_t pb[6] = {0, 0, 0, 0, 0, 0};
  if (mode == 1) { for (uint32_t s = 0; s < 2; s++) { pb[2 * s] = pb[2 * s + 1] = bits_get(b, pos, 1); pos++; } }
  else if (has_p) { for (uint32_t k = 0; k < subsets * 2; k++) { pb[k] = bits_get(b, pos, 1); pos++; } }
  for (uint32_t k = 0; k < subsets * 2; k++)
    for (uint32_t c = 0; c < 4; c++) {
      if (c >= comps) { e[k][c] = 255; continue; }
      const uint32_t nb = (c == 3) ? ap : cp;
      e[k][c] = has_p ? bc7_dq((e[k][c] << 1) | pb[k], nb + 1) : bc7_dq(e[k][c], nb);
    }
  uint32_t pat = 0, a1 = 0, a2 = 0;
  if (subsets == 2) { pat = spread_mask2(UBT(bc7_part2)[part]); a1 = UBT(bc7_anchor2)[part]; }
  else if (subsets == 3) { pat = UBT(bc7_part3)[part]; a1 = UBT(bc7_anchor3a)[part]; a2 = UBT(bc7_anchor3b)[part]; }
  for (uint32_t i = 0; i < 16; i++) {
    uint32_t nb = ib;
    if (i == 0 || (subsets >= 2 && i == a1) || (subsets == 3 && i == a2)) nb--;
    const uint32_t w = bc7_w(ib, bits_get(b, pos, nb)); pos += nb;
    const uint32_t s = (pat >> (2 * i)) & 3u;
    uint32_t v = 0;
    for (uint32_t c = 0; c < 4; c++) v |= ((c < comps) ? bc7_lerp(e[2 * s][c], e[2 * s + 1][c], w) : 255u) << (8 * c);
    out[i] = v;
  }
  return true;
}

// Texels of the BC7 block a Bc7Build describes, without packing it first.  encode_bc7_block's anchor normalisation
// (swap a subset's endpoints + p-bits, invert its selectors) does not change what the block decodes to, because the
// BC7 weight tables are symmetric (w[max - i] == 64 - w[i]); everything else it does is serialisation.  So this equals
// decode_bc7_pixels(bc7_pack(o)) -- the encoder's candidate scoring and the RDO tables use it to skip both steps.
UB_FN void decode_bc7_build(const Bc7Build& o, uint32_t* out) {
  const uint32_t m = o.mode;
  const uint32_t subsets = bc7_subsets(m), ib = bc7_index_bits(m);
  const uint32_t cp = bc7_color_prec(m), ap = bc7_alpha_prec(m);
  const bool has_p = (m == 1 || m == 3 || m == 6 || m == 7);
  uint32_t lo[3], hi[3];
#pragma unroll
  for (uint32_t k = 0; k < 3; k++) {
    uint32_t l = 0xFF000000u, h = 0xFF000000u;
    if (k < subsets) {
      const uint32_t pl = (o.pbits >> (2 * k)) & 1u, ph = (m == 1) ? pl : ((o.pbits >> (2 * k + 1)) & 1u);
      l = 0; h = 0;
#pragma unroll
      for (uint32_t c = 0; c < 4; c++) {
        uint32_t a = 255, b = 255;
        if (c < 3 || m >= 4) {
          const uint32_t nb = (c == 3) ? ap : cp;
          a = px_c(o.lo[k], c); b = px_c(o.hi[k], c);
          if (has_p) { a = bc7_dq((a << 1) | pl, nb + 1); b = bc7_dq((b << 1) | ph, nb + 1); }
          else { a = bc7_dq(a, nb); b = bc7_dq(b, nb); }
        }
        l |= a << (8 * c); h |= b << (8 * c);
      }
    }
    lo[k] = l; hi[k] = h;
  }
  uint32_t pat = 0;
  if (subsets == 2) pat = spread_mask2(UBT(bc7_part2)[o.partition]);
  else if (subsets == 3) pat = UBT(bc7_part3)[o.partition];
  const uint32_t wmask = (1u << ib) - 1u;
#pragma unroll 1
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t s = (pat >> (2 * i)) & 3u;
    const uint32_t l = s == 0 ? lo[0] : (s == 1 ? lo[1] : lo[2]), h = s == 0 ? hi[0] : (s == 1 ? hi[1] : hi[2]);
    const uint32_t w = bc7_w(ib, (uint32_t)(o.sel >> (i * ib)) & wmask);
    uint32_t v;
    if (m == 5) {
      const uint32_t wa = UBT(bc7_weights2)[(uint32_t)(o.asel >> (i * 2)) & 3u];
      uint32_t p[4];
#pragma unroll
      for (uint32_t c = 0; c < 3; c++) p[c] = bc7_lerp(px_c(l, c), px_c(h, c), w);
      p[3] = bc7_lerp(px_c(l, 3), px_c(h, 3), wa);
      const uint32_t rot = o.rotation;
      if (rot >= 1) {  // swap alpha with channel rot-1
        const uint32_t t = p[3];
        if (rot == 1) { p[3] = p[0]; p[0] = t; } else if (rot == 2) { p[3] = p[1]; p[1] = t; } else { p[3] = p[2]; p[2] = t; }
      }
      v = p[0] | (p[1] << 8) | (p[2] << 16) | (p[3] << 24);
    } else {
      v = bc7_lerp(px_c(l, 0), px_c(h, 0), w) | (bc7_lerp(px_c(l, 1), px_c(h, 1), w) << 8) | (bc7_lerp(px_c(l, 2), px_c(h, 2), w) << 16);
      v |= ((m >= 4) ? bc7_lerp(px_c(l, 3), px_c(h, 3), w) : 255u) << 24;
    }
    out[i] = v;
  }
}

}  // namespace ub
