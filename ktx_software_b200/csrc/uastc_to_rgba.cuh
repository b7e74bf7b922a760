// uastc_to_rgba.cuh -- UASTC block -> 16 RGBA8 texels for the uncompressed transcode targets (KTX_TTF_RGBA32, RGB565,
// BGR565, RGBA4444).  Restates unpack_uastc(struct -> pixels), reference
// external/basis_universal/transcoder/basisu_transcoder.cpp:15378-15514 (srgb = false, as transcode_slice calls it,
// XC:10208-10292) on top of the packed-weight unpack of uastc_unpack.cuh.
//
// Differences in form (not in result): no per-subset colour table; every texel is interpolated directly, two channels
// at a time in the 16-bit halves of a word; a solid-colour block is "both endpoints = the colour".  With l, h <= 255 and w <= 64, t = l*(64-w) + h*w fits 14 bits and
//   astc_interpolate(l, h, w) = ((l*257*(64-w) + h*257*w + 32) >> 6) >> 8 = (257*t + 32) >> 14 = (t + ((t + 32) >> 8)) >> 6,
// and the 1..5-bit weight tables are "replicate to 6 bits, +1 above 32" (checked against the generated tables by
// tests/test_oracle_cpu.py::test_weight_formula_matches_tables).
#pragma once
#include "uastc_unpack.cuh"

namespace ub {

// g_astc_weight_tables[wb][l] (XC:15376): g_bc7_weights1/2/3, g_astc_weights4/5
UB_FN uint32_t weight_unquant(uint32_t wb, uint32_t l) {
  uint32_t v;
  switch (wb) {
    case 1: v = l * 63u; break;
    case 2: v = l * 21u; break;
    case 3: v = l * 9u; break;
    case 4: v = (l << 2) | (l >> 2); break;
    default: v = (l << 1) | (l >> 4); break;
  }
  return v + (v > 32u ? 1u : 0u);
}

// two channels at once: lanes of 16 bits, each holding an 8-bit endpoint
UB_FN uint32_t interp_pair(uint32_t l2, uint32_t h2, uint32_t w) {
  const uint32_t t = l2 * (64u - w) + h2 * w;
  const uint32_t u = ((t + 0x00200020u) >> 8) & 0x00FF00FFu;
  return ((t + u) >> 6) & 0x00FF00FFu;
}

// What the interpolation needs from a block, whatever its mode.  Building it is the mode-specific part (field offsets,
// BISE layout, weight width: all constants once the mode is known); the interpolation that follows is one piece of
// code shared by every mode, so warps that hold blocks of different modes reconverge on it and the kernel's
// instruction footprint per mode stays small.
struct TexelForm {
  uint32_t l_rb[3], h_rb[3], l_ga[3], h_ga[3];   // endpoints per subset as channel pairs: rb = r | b<<16, ga = g | a<<16
  uint32_t pat;                                   // subset of texel i = (pat >> 2i) & 3
  uint32_t w0[4], w1[4];                          // unquantised weight (0..64) of texel i in byte i; w1 = second plane
  uint32_t sel;                                   // byte mask of the channel that takes the second plane (0 = one plane)
};

UB_FN void texel_form_clear(TexelForm& f) {
#pragma unroll
  for (uint32_t s = 0; s < 3; s++) f.l_rb[s] = f.h_rb[s] = f.l_ga[s] = f.h_ga[s] = 0;
#pragma unroll
  for (uint32_t k = 0; k < 4; k++) f.w0[k] = f.w1[k] = 0;
  f.pat = 0; f.sel = 0;
}

// KM >= 0: mode known by the caller (blocks binned by mode), every field offset folds to a constant.
// Returns false for an invalid block (the form then decodes to zero texels).
template <int KM, bool TS = false>
UB_FN void texel_form_from_unpacked(const UnpackedUastc& u, TexelForm& f) {   // f cleared by the caller
  const uint32_t mode = (KM >= 0) ? (uint32_t)KM : u.mode;
  if (mode == UASTC_MODE_SOLID) {   // both endpoints = the colour, weights 0: interpolates to the colour itself
    f.l_rb[0] = f.h_rb[0] = (u.solid & 0xFFu) | (((u.solid >> 16) & 0xFFu) << 16);
    f.l_ga[0] = f.h_ga[0] = ((u.solid >> 8) & 0xFFu) | ((u.solid >> 24) << 16);
    return;
  }
  const uint32_t d = (KM >= 0) ? mode_desc_c(KM) : u.desc;
  const uint32_t subsets = md_subsets(d), comps = md_comps(d), er = md_endpoint_range(d), wb = md_weight_bits(d);
  const uint32_t planes = 1 + md_dual_plane(d);
#pragma unroll
  for (uint32_t s = 0; s < 3; s++) {
    if (s < subsets) {
      if (comps == 2) {   // luminance + alpha (XC:15400-15411)
        const uint32_t ll = astc_unquant_s<TS>(er, u.ep[s * 4 + 0]), lh = astc_unquant_s<TS>(er, u.ep[s * 4 + 1]);
        const uint32_t al = astc_unquant_s<TS>(er, u.ep[s * 4 + 2]), ah = astc_unquant_s<TS>(er, u.ep[s * 4 + 3]);
        f.l_rb[s] = ll * 0x00010001u; f.h_rb[s] = lh * 0x00010001u;
        f.l_ga[s] = ll | (al << 16);  f.h_ga[s] = lh | (ah << 16);
      } else {
        uint32_t lo[4], hi[4];
#pragma unroll
        for (uint32_t k = 0; k < 4; k++) {
          lo[k] = 255; hi[k] = 255;
          if (k < comps) { lo[k] = astc_unquant_s<TS>(er, u.ep[s * comps * 2 + k * 2]); hi[k] = astc_unquant_s<TS>(er, u.ep[s * comps * 2 + k * 2 + 1]); }
        }
        f.l_rb[s] = lo[0] | (lo[2] << 16); f.h_rb[s] = hi[0] | (hi[2] << 16);
        f.l_ga[s] = lo[1] | (lo[3] << 16); f.h_ga[s] = hi[1] | (hi[3] << 16);
      }
    }
  }
  f.pat = u.pat;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) {
    f.w0[i >> 2] |= weight_unquant(wb, bits_get(u.w, (i * planes) * wb, wb)) << (8 * (i & 3u));
    if (planes == 2) f.w1[i >> 2] |= weight_unquant(wb, bits_get(u.w, (i * planes + 1) * wb, wb)) << (8 * (i & 3u));
  }
  if (planes == 2) f.sel = 255u << (8 * u.ccs);   // channel ccs takes the texel's second weight (XC:15490-15507)
}

template <int KM, bool TS = false>
UB_FN bool texel_form_t(const Bits128& src, TexelForm& f) {
  texel_form_clear(f);
  UnpackedUastc u;
  if (!unpack_uastc_t<KM, TS>(src, u, false)) return false;
  texel_form_from_unpacked<KM, TS>(u, f);
  return true;
}

// The shared part: 16 texels (r | g<<8 | b<<16 | a<<24), raster order.
UB_FN void texels_from_form(const TexelForm& f, uint32_t* px) {
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t s = (f.pat >> (2 * i)) & 3u;
    const uint32_t a = s == 0 ? f.l_rb[0] : (s == 1 ? f.l_rb[1] : f.l_rb[2]);
    const uint32_t b = s == 0 ? f.h_rb[0] : (s == 1 ? f.h_rb[1] : f.h_rb[2]);
    const uint32_t c = s == 0 ? f.l_ga[0] : (s == 1 ? f.l_ga[1] : f.l_ga[2]);
    const uint32_t e = s == 0 ? f.h_ga[0] : (s == 1 ? f.h_ga[1] : f.h_ga[2]);
    const uint32_t w = (f.w0[i >> 2] >> (8 * (i & 3u))) & 255u;
    px[i] = interp_pair(a, b, w) | (interp_pair(c, e, w) << 8);
  }
  if (f.sel) {
#pragma unroll
    for (uint32_t i = 0; i < 16; i++) {   // dual plane: one subset
      const uint32_t w = (f.w1[i >> 2] >> (8 * (i & 3u))) & 255u;
      const uint32_t p1 = interp_pair(f.l_rb[0], f.h_rb[0], w) | (interp_pair(f.l_ga[0], f.h_ga[0], w) << 8);
      px[i] = (px[i] & ~f.sel) | (p1 & f.sel);
    }
  }
}

// Block -> 16 texels.  Returns false for an invalid block (texels zeroed).
template <int KM>
UB_FN bool decode_block_rgba_t(const Bits128& src, uint32_t* px) {
  TexelForm f;
  const bool ok = texel_form_t<KM>(src, f);
  texels_from_form(f, px);
  return ok;
}

UB_FN bool decode_block_rgba(const Bits128& src, uint32_t* px) { return decode_block_rgba_t<-1>(src, px); }

// 8-bit channel -> q+1 levels, round to nearest (mul_8, XC:269)
UB_FN uint32_t mul_8(uint32_t v, uint32_t q) { v = v * q + 128u; return ((v + (v >> 8)) >> 8) & 255u; }

// RGBA8 texel -> the 16-bit targets of transcode_slice (XC:10230-10286); kind: 14 = RGB565, 15 = BGR565, 16 = RGBA4444
UB_FN uint32_t pack_texel16(uint32_t p, uint32_t kind) {
  const uint32_t r = p & 255u, g = (p >> 8) & 255u, b = (p >> 16) & 255u, a = p >> 24;
  if (kind == 14) return (mul_8(r, 31) << 11) | (mul_8(g, 63) << 5) | mul_8(b, 31);
  if (kind == 15) return (mul_8(b, 31) << 11) | (mul_8(g, 63) << 5) | mul_8(r, 31);
  return (mul_8(r, 15) << 12) | (mul_8(g, 15) << 8) | (mul_8(b, 15) << 4) | mul_8(a, 15);
}

}  // namespace ub
