// uastc_to_bcn.cuh -- the hint-driven UASTC transcode targets: KTX_TTF_ETC1_RGB, ETC2_RGBA, BC1_RGB, BC3_RGBA, BC4_R,
// BC5_RG.  Restates, from external/basis_universal/transcoder/basisu_transcoder.cpp:
//   transcode_uastc_to_bc1 :18365 (hint0 :18237, hint1 :18335, encode_bc1 :17682, encode_bc1_solid_block :17634),
//   transcode_uastc_to_bc3 :18404, transcode_uastc_to_bc4 :18442, transcode_uastc_to_bc5 :18470, encode_bc4 :17372,
//   transcode_uastc_to_etc1 :16355 (apply_etc1_bias :16182, etc1_determine_selectors :16253),
//   transcode_uastc_to_etc2_rgba :17341 (transcode_uastc_to_etc2_eac_a8 :17264).
//
// Form: like the uncompressed targets, a block is first brought into a mode-independent form by mode-specialised
// code (TexelForm + the hint bits + plane-0 weights already translated to BC1 selectors); everything after that -- the
// texel interpolation, BC4 / BC1 / ETC1 / EAC packing -- is one piece of code shared by all 19 modes.  A solid-colour
// block needs no special case for the BCn targets: its 16 equal texels take the solid paths of encode_bc1 / encode_bc4,
// which produce what encode_bc1_solid_block / write_bc4_solid_block do.
#pragma once
#include "uastc_to_rgba.cuh"
#include "uastc_hints.cuh"

namespace ub {

struct HintedForm {
  TexelForm f;
  uint32_t mode;       // 0..18 (8 for an invalid block)
  uint32_t hints;      // raw hint bits of the block (unpack_uastc :14947-15007)
  uint32_t bc1_sels;   // plane-0 weights through s_uastc{1..5}_to_bc1 (XC:17365-17369), texel i at bits 2i
  uint32_t solid;      // r | g<<8 | b<<16 | a<<24 (mode 8)
};

// Mode-specific part.  Returns false for an invalid block (the form then yields a zero output block).
template <int KM, bool TS = false>
UB_FN bool hinted_form_t(const Bits128& src, HintedForm& h) {
  texel_form_clear(h.f);
  h.mode = UASTC_MODE_SOLID; h.hints = 0; h.bc1_sels = 0; h.solid = 0;
  UnpackedUastc u;
  if (!unpack_uastc_t<KM, TS>(src, u, true)) return false;
  texel_form_from_unpacked<KM, TS>(u, h.f);
  const uint32_t mode = (KM >= 0) ? (uint32_t)KM : u.mode;
  h.mode = mode; h.hints = u.hints; h.solid = u.solid;
  if (mode != UASTC_MODE_SOLID) {
    const uint32_t d = (KM >= 0) ? mode_desc_c(KM) : u.desc;
    const uint32_t wb = md_weight_bits(d), planes = 1 + md_dual_plane(d);
    uint32_t s = 0;
#pragma unroll
    for (uint32_t i = 0; i < 16; i++) s |= uastc_to_bc1_sel(wb, bits_get(u.w, (i * planes) * wb, wb)) << (2 * i);
    h.bc1_sels = s;
  }
  return true;
}

// The hint fields of a block (XC:14960-15007).  Solid blocks: diff, inten0, selector, r, g, b; others: the rest.
struct HintFields {
  uint32_t bc1_hint0, bc1_hint1, etc1_flip, etc1_diff, etc1_inten0, etc1_inten1, etc1_bias, etc2_hints;
  uint32_t etc1_selector, etc1_r, etc1_g, etc1_b;
  uint32_t has_bias, has_alpha;
};

UB_FN HintFields parse_hints(uint32_t mode, uint32_t hints) {
  HintFields o;
  o.bc1_hint0 = o.bc1_hint1 = o.etc1_flip = o.etc1_diff = o.etc1_inten0 = o.etc1_inten1 = o.etc1_bias = o.etc2_hints = 0;
  o.etc1_selector = o.etc1_r = o.etc1_g = o.etc1_b = 0;
  o.has_bias = 0; o.has_alpha = (mode >= 8 && mode <= 17) ? 1u : 0u;   // g_uastc_mode_has_alpha (XC:14060)
  uint32_t p = 0;
  if (mode == UASTC_MODE_SOLID) {
    o.etc1_diff = hints & 1u;
    o.etc1_inten0 = (hints >> 1) & 7u;
    o.etc1_selector = (hints >> 4) & 3u;
    o.etc1_r = (hints >> 6) & 31u; o.etc1_g = (hints >> 11) & 31u; o.etc1_b = (hints >> 16) & 31u;
    return o;
  }
  o.has_bias = UBT(uastc_mode_has_etc1_bias)[mode];
  if (UBT(uastc_mode_has_bc1_hint0)[mode]) { o.bc1_hint0 = (hints >> p) & 1u; p++; }
  if (UBT(uastc_mode_has_bc1_hint1)[mode]) { o.bc1_hint1 = (hints >> p) & 1u; p++; }
  o.etc1_flip = (hints >> p) & 1u; p++;
  o.etc1_diff = (hints >> p) & 1u; p++;
  o.etc1_inten0 = (hints >> p) & 7u; p += 3;
  o.etc1_inten1 = (hints >> p) & 7u; p += 3;
  if (o.has_bias) { o.etc1_bias = (hints >> p) & 31u; p += 5; }
  if (o.has_alpha) o.etc2_hints = (hints >> p) & 255u;
  return o;
}

// One channel of the 16 texels (what unpack_uastc :15378-15514 yields for it), without interpolating the other three:
// BC4 / BC5 need one or two channels only.
UB_FN void channel_from_form(const TexelForm& f, uint32_t chan, uint32_t* v) {
  const uint32_t sh = (chan & 2u) ? 16u : 0u;   // r: rb low half, g: ga low, b: rb high, a: ga high
  const bool ga = (chan & 1u) != 0;
  uint32_t l[3], h[3];
#pragma unroll
  for (uint32_t s = 0; s < 3; s++) { l[s] = ((ga ? f.l_ga[s] : f.l_rb[s]) >> sh) & 255u; h[s] = ((ga ? f.h_ga[s] : f.h_rb[s]) >> sh) & 255u; }
  const bool second = ((f.sel >> (8 * chan)) & 1u) != 0;   // dual plane and this channel takes the second weight
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t s = (f.pat >> (2 * i)) & 3u;
    const uint32_t a = s == 0 ? l[0] : (s == 1 ? l[1] : l[2]), b = s == 0 ? h[0] : (s == 1 ? h[1] : h[2]);
    const uint32_t w = ((second ? f.w1[i >> 2] : f.w0[i >> 2]) >> (8 * (i & 3u))) & 255u;
    const uint32_t t = a * (64u - w) + b * w;
    v[i] = (t + ((t + 32u) >> 8)) >> 6;   // astc_interpolate, see uastc_to_rgba.cuh
  }
}

// encode_bc4 (XC:17372-17490) of 16 channel values -> the 8-byte block, little-endian.
UB_FN uint64_t bc4_encode_values(const uint32_t* val) {
  uint32_t mn = 255u, mx = 0u;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) { mn = minu(mn, val[i]); mx = maxu(mx, val[i]); }
  uint64_t out = (uint64_t)mx | ((uint64_t)mn << 8);
  if (mx == mn) return out;
  const int delta = (int)(mx - mn), bias = 4 - (int)mn * 14;
  uint64_t bits = 0;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) {
    const int v = (int)val[i] * 14 + bias;
    const uint32_t k = (uint32_t)(v >= delta * 13) + (v >= delta * 11) + (v >= delta * 9) + (v >= delta * 7) + (v >= delta * 5) + (v >= delta * 3) + (v >= delta);
    bits |= (uint64_t)((0x02345671u >> (4 * k)) & 7u) << (3 * i);   // s_tran = {1,7,6,5,4,3,2,0}
  }
  return out | (bits << 16);
}
UB_FN uint64_t bc4_encode(const uint32_t* px, uint32_t chan) {   // channel `chan` of 16 RGBA texels
  uint32_t v[16];
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) v[i] = ch(px[i], chan);
  return bc4_encode_values(v);
}
UB_FN uint64_t bc4_from_form(const TexelForm& f, uint32_t chan) {
  uint32_t v[16];
  channel_from_form(f, chan, v);
  return bc4_encode_values(v);
}

UB_FN uint64_t bc1_bytes(const Bc1Block& b) { return (uint64_t)(b.lo16 & 0xFFFFu) | ((uint64_t)(b.hi16 & 0xFFFFu) << 16) | ((uint64_t)b.sels << 32); }

// transcode_uastc_to_bc1_hint0 (XC:18237): endpoints of subset 0 as the BC1 colours, weights as selectors.
UB_FN Bc1Block bc1_from_hint0(const HintedForm& h) {
  const uint32_t l[3] = {h.f.l_rb[0] & 255u, h.f.l_ga[0] & 255u, (h.f.l_rb[0] >> 16) & 255u};
  const uint32_t e[3] = {h.f.h_rb[0] & 255u, h.f.h_ga[0] & 255u, (h.f.h_rb[0] >> 16) & 255u};
  Bc1Block b;
  b.lo16 = ((l[0] * 31 + 127) / 255) << 11 | ((l[1] * 63 + 127) / 255) << 5 | ((l[2] * 31 + 127) / 255);
  b.hi16 = ((e[0] * 31 + 127) / 255) << 11 | ((e[1] * 63 + 127) / 255) << 5 | ((e[2] * 31 + 127) / 255);
  if (b.lo16 == b.hi16) {
    uint32_t mask = 0;
    if (b.hi16 > 0) b.hi16--; else { b.hi16 = 0; b.lo16 = 1; mask = 0x55; }
    b.sels = mask * 0x01010101u;
  } else {
    b.sels = h.bc1_sels;
    if (b.lo16 < b.hi16) { const uint32_t t = b.lo16; b.lo16 = b.hi16; b.hi16 = t; b.sels ^= 0x55555555u; }
  }
  return b;
}

// The BC1 colour block of transcode_uastc_to_bc1 / _bc3 (XC:18377-18392, :18428-18438)
UB_FN uint64_t bc1_from_form(const HintedForm& h, const HintFields& hf, const uint32_t* px, bool high_quality) {
  if (!high_quality && hf.bc1_hint0) return bc1_bytes(bc1_from_hint0(h));
  uint8_t sels[16];
  const uint32_t passes = high_quality ? 2u : 1u;
  if (hf.bc1_hint1) {
    // transcode_uastc_to_bc1_hint1: BC1 selector order -> linear order {0,3,1,2} (s_sel_tran, XC:17697)
#pragma unroll
    for (uint32_t i = 0; i < 16; i++) sels[i] = (uint8_t)((0x9Cu >> (2 * ((h.bc1_sels >> (2 * i)) & 3u))) & 3u);
    return bc1_bytes(encode_bc1(px, true, sels, passes));
  }
  return bc1_bytes(encode_bc1(px, false, sels, passes));
}

// ---- ETC1 (XC:16355-16441)
// the four colours of a sub-block: decoder_etc_block::get_block_colors (XC:784), luma-weighted as etc1_determine_selectors does
UB_FN void etc1_subblock_luma(const uint32_t* base8, uint32_t inten, uint32_t* y) {
#pragma unroll
  for (uint32_t k = 0; k < 4; k++) {
    const int m = UBT(etc1_inten)[inten * 4 + k];
    y[k] = (uint32_t)clamp255((int)base8[0] + m) * 54u + (uint32_t)clamp255((int)base8[1] + m) * 183u + (uint32_t)clamp255((int)base8[2] + m) * 19u;
  }
}

UB_FN uint64_t etc1_from_form(const HintedForm& h, const HintFields& hf, const uint32_t* px) {
  uint32_t by[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (h.mode == UASTC_MODE_SOLID) {
    by[3] = ((hf.etc1_diff << 1) | (hf.etc1_inten0 << 5) | (hf.etc1_inten0 << 2)) & 255u;
    if (hf.etc1_diff) { by[0] = (hf.etc1_r << 3) & 255u; by[1] = (hf.etc1_g << 3) & 255u; by[2] = (hf.etc1_b << 3) & 255u; }
    else { by[0] = (hf.etc1_r | (hf.etc1_r << 4)) & 255u; by[1] = (hf.etc1_g | (hf.etc1_g << 4)) & 255u; by[2] = (hf.etc1_b | (hf.etc1_b << 4)) & 255u; }
    // s_etc1_solid_selectors (XC:16313)
    const uint32_t s = hf.etc1_selector;
    by[4] = by[5] = (s < 2) ? 255u : 0u;
    by[6] = by[7] = (s == 0 || s == 3) ? 255u : 0u;
  } else {
    const uint32_t flip = hf.etc1_flip, diff = hf.etc1_diff;
    by[3] = (flip | (diff << 1) | (hf.etc1_inten0 << 5) | (hf.etc1_inten1 << 2)) & 255u;
    const uint32_t limit = diff ? 31u : 15u;
    int bc[2][3];
#pragma unroll
    for (uint32_t subset = 0; subset < 2; subset++) {
      uint32_t avg[3] = {0, 0, 0};
#pragma unroll
      for (uint32_t j = 0; j < 8; j++) {
        const uint32_t p = px[etc1_texel(flip, subset, j)];
        avg[0] += ch(p, 0); avg[1] += ch(p, 1); avg[2] += ch(p, 2);
      }
#pragma unroll
      for (uint32_t c = 0; c < 3; c++) {
        int v = (int)((avg[c] * limit + 1020u) / (8u * 255u));
        if (hf.has_bias) v = etc1_bias_comp(v, hf.etc1_bias, (int)limit, subset, c);
        bc[subset][c] = v;
      }
    }
    uint32_t base8[2][3];
    if (diff) {
#pragma unroll
      for (uint32_t c = 0; c < 3; c++) {
        int d = clampi(bc[1][c] - bc[0][c], -4, 3);
        // what the decoder sees for sub-block 1: base + delta, clamped to 5 bits (unpack_color5, XC:907)
        const int c1 = clampi(bc[0][c] + d, 0, 31);
        base8[0][c] = (uint32_t)((bc[0][c] << 3) | (bc[0][c] >> 2));
        base8[1][c] = (uint32_t)((c1 << 3) | (c1 >> 2));
        if (d < 0) d += 8;
        by[c] = (uint32_t)((bc[0][c] << 3) | d) & 255u;
      }
    } else {
#pragma unroll
      for (uint32_t c = 0; c < 3; c++) {
        by[c] = (uint32_t)(bc[1][c] | (bc[0][c] << 4)) & 255u;
        base8[0][c] = (uint32_t)((bc[0][c] << 4) | bc[0][c]);
        base8[1][c] = (uint32_t)((bc[1][c] << 4) | bc[1][c]);
      }
    }
    // etc1_determine_selectors (XC:16253): ETC1 bit index of texel (x, y) = x*4 + y
    uint32_t lmask = 0, hmask = 0;
#pragma unroll
    for (uint32_t subblock = 0; subblock < 2; subblock++) {
      uint32_t y4[4];
      etc1_subblock_luma(base8[subblock], subblock ? hf.etc1_inten1 : hf.etc1_inten0, y4);
      const uint32_t y01 = y4[0] + y4[1], y12 = y4[1] + y4[2], y23 = y4[2] + y4[3];
#pragma unroll
      for (uint32_t j = 0; j < 8; j++) {
        const uint32_t ti = etc1_texel(flip, subblock, j), x = ti & 3u, yy = ti >> 2;
        const uint32_t p = px[ti];
        const uint32_t l = ch(p, 0) * 108u + ch(p, 1) * 366u + ch(p, 2) * 38u;
        const uint32_t t = (0xE1u >> (2 * ((l < y01) + (l < y12) + (l < y23)))) & 3u;   // s_tran = {1,0,2,3}
        const uint32_t ofs = x * 4 + yy;
        lmask |= (t & 1u) << ofs;
        hmask |= (t >> 1) << ofs;
      }
    }
    by[7] = lmask & 255u; by[6] = (lmask >> 8) & 255u; by[5] = hmask & 255u; by[4] = (hmask >> 8) & 255u;
  }
  uint64_t o = 0;
#pragma unroll
  for (uint32_t i = 0; i < 8; i++) o |= (uint64_t)by[i] << (8 * i);
  return o;
}

// ---- ETC1 from ONE channel (transcode_uastc_to_etc1(blk, dst, channel), XC:16560-17260; transcode_slice passes channel 3 when the
// caller asks for alpha data in an opaque format, :10142-10147).  There are no hints for this: flip estimate, base grey levels
// and the intensity tables are searched here.  val = the channel's 16 values, raster order.
#include "uastc_etc1y_tables.inc"
UB_FN uint64_t etc1_y_bytes(const uint32_t* by) {
  uint64_t o = 0;
#pragma unroll
  for (uint32_t i = 0; i < 8; i++) o |= (uint64_t)(by[i] & 255u) << (8 * i);
  return o;
}
UB_FN uint64_t etc1_y_solid(uint32_t y) {
  const uint32_t enc = UBT(etc1y_solid)[y], base = enc & 31u, sel = (enc >> 5) & 3u, table = enc >> 7;
  uint32_t by[8];
  by[0] = by[1] = by[2] = base << 3;
  by[3] = 2u | (table << 5) | (table << 2);
  by[4] = by[5] = (sel < 2) ? 255u : 0u;                 // s_etc1_solid_selectors (XC:16323)
  by[6] = by[7] = (sel == 0 || sel == 3) ? 255u : 0u;
  return etc1_y_bytes(by);
}
UB_BIG uint64_t etc1_y_from_values(const uint32_t* val) {
  // pack_etc1_y_estimate_flipped (XC:16469): quadrant sums -> half averages -> which split leaves the greyer halves
  int q[2][2] = {{0, 0}, {0, 0}};   // [x half][y half]
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) q[(i & 3u) >> 1][i >> 3] += (int)val[i];
  const int upper = (q[0][0] + q[1][0] + 4) / 8, lower = (q[0][1] + q[1][1] + 4) / 8, left = (q[0][0] + q[0][1] + 4) / 8, right = (q[1][0] + q[1][1] + 4) / 8;
  int ud = 0, ld = 0, lfd = 0, rd = 0;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) {
    const int v = (int)val[i];
    const int dv = v - ((i >> 3) ? lower : upper), dh = v - (((i & 3u) >> 1) ? right : left);
    if (i >> 3) ld += dv * dv; else ud += dv * dv;
    if ((i & 3u) >> 1) rd += dh * dh; else lfd += dh * dh;
  }
  const uint32_t flip = (ud + ld) < (lfd + rd) ? 1u : 0u;
  uint32_t low[2] = {255u, 255u}, high[2] = {0u, 0u};
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t sb = flip ? (i >> 3) : ((i & 3u) >> 1);
    low[sb] = minu(low[sb], val[i]); high[sb] = maxu(high[sb], val[i]);
  }
  const uint32_t range[2] = {high[0] - low[0], high[1] - low[1]};
  uint32_t by[8] = {0, 0, 0, flip, 0, 0, 0, 0};
  uint32_t lmask = 0, hmask = 0;
  // a sub-block whose values span <= 3 levels: base, table and selectors straight from the run tables (individual mode)
  auto run_subblock = [&](uint32_t sb) {
    const uint32_t enc = (range[sb] == 0) ? UBT(etc1y_run1)[low[sb]] : ((range[sb] < 2) ? UBT(etc1y_run2)[low[sb]] : UBT(etc1y_run4)[low[sb]]);
    const uint32_t table = enc & 7u, base = (enc >> 3) & 31u;
    by[3] |= table << (sb ? 2 : 5);
    const uint32_t keep = ~(0xFu << (sb ? 0 : 4)), sv = base << (sb ? 0 : 4);
    by[0] = (by[0] & keep) | sv; by[1] = (by[1] & keep) | sv; by[2] = (by[2] & keep) | sv;
#pragma unroll
    for (uint32_t i = 0; i < 16; i++) {
      const uint32_t mine = flip ? (i >> 3) : ((i & 3u) >> 1);
      if (mine == sb) {
        const uint32_t s = (enc >> (8 + 2 * (val[i] - low[sb]))) & 3u;
        const uint32_t t = (0x4Bu >> (2 * s)) & 3u;                 // g_selector_index_to_etc1 = {3, 2, 0, 1}
        const uint32_t ofs = (i & 3u) * 4 + (i >> 2);
        lmask |= (t & 1u) << ofs; hmask |= (t >> 1) << ofs;
      }
    }
  };
  if (range[0] <= 3 && range[1] <= 3) {
    run_subblock(0); run_subblock(1);
  } else {
    const int a0 = flip ? upper : left, a1 = flip ? lower : right;
    uint32_t y0 = (uint32_t)(a0 * 31 + 127) / 255u, y1 = (uint32_t)(a1 * 31 + 127) / 255u;
    bool diff = true;
    int dy = (int)y1 - (int)y0;
    if (dy < -4 || dy > 3) {
      diff = false;
      y0 = (uint32_t)(a0 * 15 + 127) / 255u; y1 = (uint32_t)(a1 * 15 + 127) / 255u;
      by[0] = by[1] = by[2] = y1 | (y0 << 4);
    } else {
      y1 = (uint32_t)((int)y0 + dy);
      if (dy < 0) dy += 8;
      by[0] = by[1] = by[2] = (y0 << 3) | (uint32_t)dy;
      by[3] |= 2u;
    }
    const uint32_t base_y[2] = {diff ? ((y0 << 3) | (y0 >> 2)) : ((y0 << 4) | y0), diff ? ((y1 << 3) | (y1 >> 2)) : ((y1 << 4) | y1)};
#pragma unroll 1
    for (uint32_t sb = 0; sb < 2; sb++) {
      if (!diff && range[sb] <= 3) { run_subblock(sb); continue; }
      const int base = (int)base_y[sb];
      int pos = (int)high[sb] - base, neg = base - (int)low[sb];
      if (pos < 0) pos = -pos;
      if (neg < 0) neg = -neg;
      const uint32_t er = (uint32_t)(pos > neg ? pos : neg);
      // only the tables that can be useful for this range (XC:17019-17053)
      uint32_t mask = 0xFFu;
      if (er <= 51) {
        mask = 0xFu;
        if (er > 22) mask &= ~1u;
        if (er < 4 || er > 39) mask &= ~2u;
        if (er < 9) mask &= ~4u;
        if (er < 12) mask &= ~8u;
      } else {
        mask &= ~3u;
        if (er > 60) mask &= ~4u;
        if (er > 89) mask &= ~8u;
        if (er > 120) mask &= ~16u;
        if (er > 136) mask &= ~32u;
        if (er > 174) mask &= ~64u;
      }
      const int low_limit = -base, high_limit = 255 - base;
      uint32_t best_err = 0xFFFFFFFFu, best_inten = 0, best_l = 0, best_h = 0;
#pragma unroll 1
      for (uint32_t inten = 0; inten < 8; inten++) {
        if (!((mask >> inten) & 1u)) continue;
        int t0 = UBT(etc1_inten)[inten * 4 + 0], t1 = UBT(etc1_inten)[inten * 4 + 1], t2 = UBT(etc1_inten)[inten * 4 + 2], t3 = UBT(etc1_inten)[inten * 4 + 3];
        t0 = t0 > low_limit ? t0 : low_limit; t1 = t1 > low_limit ? t1 : low_limit;
        t2 = t2 < high_limit ? t2 : high_limit; t3 = t3 < high_limit ? t3 : high_limit;
        const int th01 = t0 + t1, th12 = t1 + t2, th23 = t2 + t3;
        uint32_t total = 0, lm = 0, hm = 0;
#pragma unroll
        for (uint32_t i = 0; i < 16; i++) {
          const uint32_t mine = flip ? (i >> 3) : ((i & 3u) >> 1);
          if (mine == sb) {
            const int delta = (int)val[i] - base, d2 = delta * 2;
            const uint32_t k = (uint32_t)(d2 < th01) + (d2 < th12) + (d2 < th23);
            const uint32_t c = (0xE1u >> (2 * k)) & 3u;                       // s_table = {1, 0, 2, 3}
            const int tv = k == 0 ? t3 : (k == 1 ? t2 : (k == 2 ? t1 : t0));  // tv[] = {t2, t3, t1, t0} indexed by c
            const int e = tv - delta;
            total += (uint32_t)(e * e);
            const uint32_t ofs = (i & 3u) * 4 + (i >> 2);
            lm |= (c & 1u) << ofs; hm |= (c >> 1) << ofs;
          }
        }
        if (total < best_err) { best_err = total; best_inten = inten; best_l = lm; best_h = hm; }
      }
      by[3] |= best_inten << (sb ? 2 : 5);
      lmask |= best_l; hmask |= best_h;
    }
  }
  by[7] = lmask & 255u; by[6] = (lmask >> 8) & 255u; by[5] = hmask & 255u; by[4] = (hmask >> 8) & 255u;
  return etc1_y_bytes(by);
}

// ---- ETC2 EAC A8 (XC:17264-17339)
UB_FN uint64_t eac_a8_bytes(uint32_t base, uint32_t table, uint32_t mul, uint64_t sels48) {
  uint64_t o = (uint64_t)(base & 255u) | ((uint64_t)((table & 15u) | ((mul & 15u) << 4)) << 8);
#pragma unroll
  for (uint32_t k = 0; k < 6; k++) o |= ((sels48 >> (40 - 8 * k)) & 255ull) << (8 * (2 + k));   // set_selector_bits: big-endian
  return o;
}

UB_FN uint64_t eac_a8_from_form(const HintedForm& h, const HintFields& hf, const uint32_t* px) {
  const uint64_t SEL4 = 0x924924924924ull;   // g_etc2_eac_a8_sel4 (XC:4781)
  if (!hf.has_alpha || h.mode == UASTC_MODE_SOLID) return eac_a8_bytes(h.mode == UASTC_MODE_SOLID ? (h.solid >> 24) : 255u, 13, 1, SEL4);
  uint32_t mn = 255u, mx = 0u;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) { const uint32_t a = px[i] >> 24; mn = minu(mn, a); mx = maxu(mx, a); }
  if (mn == mx) return eac_a8_bytes(mn, 13, 1, SEL4);
  const uint32_t table = hf.etc2_hints & 15u;
  const int mul = (int)(hf.etc2_hints >> 4);
  const int8_t* T = UBT(etc2_eac) + table * 8;
  const float range = (float)((int)T[7] - (int)T[3]);
  const float t = f_div((float)(0 - (int)T[3]), range);
  const int center = (int)roundf(f_add((float)mn, f_mul(f_sub((float)mx, (float)mn), t)));   // basisu::lerp = a + (b - a) * t
  int vals[8];
#pragma unroll
  for (uint32_t j = 0; j < 8; j++) vals[j] = clamp255(center + (int)T[j] * mul);
  uint64_t sels = 0;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) {
    const int a = (int)(px[(i & 3u) * 4 + (i >> 2)] >> 24);   // block_pixels[i & 3][i >> 2]: column-major
    uint32_t best = 0xFFFFFFFFu;
#pragma unroll
    for (uint32_t j = 0; j < 8; j++) {
      int e = vals[j] - a;
      if (e < 0) e = -e;
      best = minu(best, ((uint32_t)e << 3) | j);
    }
    sels |= (uint64_t)(best & 7u) << (45 - i * 3);
  }
  return eac_a8_bytes((uint32_t)center, table, (uint32_t)mul, sels);
}

// ---- ETC2 EAC R11 / RG11 (XC:18792-18850): pack_eac (:18515, four tables) or pack_eac_high_quality (:18672, all sixteen)
// on one channel's 16 values; a solid block has multiplier 0 (pack_eac_solid_block :18503), unlike the A8 path's 1.
UB_FN uint64_t eac_pack_values(const uint32_t* val, bool high_quality) {
  const uint64_t SEL4 = 0x924924924924ull;
  uint32_t mn = 255u, mx = 0u;
#pragma unroll
  for (uint32_t i = 0; i < 16; i++) { mn = minu(mn, val[i]); mx = maxu(mx, val[i]); }
  if (mn == mx) return eac_a8_bytes(mn, 13, 0, SEL4);
  const uint32_t arange = mx - mn;
  if (arange <= 5u) {   // table 13 is lossless
    const int base = clamp255((int)mx - 2);
    uint64_t sels = 0;
#pragma unroll
    for (uint32_t i = 0; i < 16; i++) {
      const uint32_t sel = (uint32_t)((int)val[i] - (base - 3));                    // 0..5
      sels |= (uint64_t)((0x654012u >> (4 * sel)) & 7u) << (45 - 3 * ((i & 3u) * 4 + (i >> 2)));   // s_sels = {2,1,0,4,5,6}
    }
    return eac_a8_bytes((uint32_t)base, 13, 1, sels);
  }
  uint32_t best_err = 0xFFFFFFFFu, b_base = 0, b_mul = 0, b_table = 0;
  uint64_t b_sels = 0;
  const uint32_t n_tables = high_quality ? 16u : 4u;
  for (uint32_t k = 0; k < n_tables; k++) {
    const uint32_t table = high_quality ? k : ((0xDB82u >> (4 * k)) & 15u);   // {2, 8, 11, 13}
    const int8_t* T = UBT(etc2_eac) + table * 8;
    const float range = (float)((int)T[7] - (int)T[3]);
    const float t = f_div((float)(0 - (int)T[3]), range);
    const int base = clamp255((int)roundf(f_add((float)mn, f_mul(f_sub((float)mx, (float)mn), t))));
    const int mul = clampi((int)roundf(f_div((float)arange, range)), 1, 15);
    uint32_t err = 0;
    uint64_t sels = 0;
#pragma unroll
    for (uint32_t i = 0; i < 16; i++) {
      const int a = (int)val[i];
      const bool clamped = high_quality || a < 7 || a > 255 - 7;   // pack_eac clamps the table values only near the ends (:18592-18636)
      uint32_t l = 0xFFFFFFFFu;
#pragma unroll
      for (uint32_t sidx = 0; sidx < 8; sidx++) {
        int v = mul * (int)T[sidx] + base;
        if (clamped) v = clamp255(v);
        int e = v - a;
        if (e < 0) e = -e;
        l = minu(l, ((uint32_t)e << 3) | sidx);
      }
      sels |= (uint64_t)(l & 7u) << (45 - 3 * ((i & 3u) * 4 + (i >> 2)));
      err += (l >> 3) * (l >> 3);
    }
    if (err < best_err) { best_err = err; b_base = (uint32_t)base; b_mul = (uint32_t)mul; b_table = table; b_sels = sels; }
  }
  return eac_a8_bytes(b_base, b_table, b_mul, b_sels);
}

// one EAC block of transcode_uastc_to_etc2_eac_r11 / _rg11 for channel `chan`
UB_FN uint64_t eac_r11_from_form(const HintedForm& h, const HintFields& hf, uint32_t chan, bool high_quality) {
  if (h.mode == UASTC_MODE_SOLID) return eac_a8_bytes((h.solid >> (8 * chan)) & 255u, 13, 0, 0x924924924924ull);
  if (chan == 3) {
    uint32_t px[16];
    texels_from_form(h.f, px);
    return eac_a8_from_form(h, hf, px);
  }
  uint32_t v[16];
  channel_from_form(h.f, chan, v);
  return eac_pack_values(v, high_quality);
}

// Output bytes per block of a hinted target (0 = not one)
UB_FN uint32_t hinted_target_bytes(uint32_t kind) { return (kind == 0 || kind == 2 || kind == 4 || kind == 6 || kind == 8) ? 8u : (kind == 1 || kind == 3 || kind == 5 || kind == 7) ? 16u : 0u; }

// The shared part.  kind = ktx_transcode_fmt_e for 0 ETC1_RGB, 1 ETC2_RGBA, 2 BC1_RGB, 3 BC3_RGBA, 4 BC4_R, 5 BC5_RG, and
// 6 = ETC2_EAC_R11, 7 = ETC2_EAC_RG11 (KTX_TTF 20 / 21), 8 = ETC1_RGB from the single channel chan0; chan0 / chan1: the channels BC4 / BC5 / R11 / RG11 take
// (transcode_image, XC:10417-10438, :10511-10527).  8-byte targets fill d.lo only.
UB_FN void hinted_target_from_form(const HintedForm& h, uint32_t kind, bool high_quality, uint32_t chan0, uint32_t chan1, Bits128& d) {
  d.lo = 0; d.hi = 0;
  if (kind == 8) {   // ETC1 from channel chan0 (alpha data to an opaque format)
    if (h.mode == UASTC_MODE_SOLID) { d.lo = etc1_y_solid((h.solid >> (8 * chan0)) & 255u); return; }
    uint32_t v[16];
    channel_from_form(h.f, chan0, v);
    d.lo = etc1_y_from_values(v);
    return;
  }
  if (kind == 4) { d.lo = bc4_from_form(h.f, chan0); return; }
  if (kind == 5) { d.lo = bc4_from_form(h.f, chan0); d.hi = bc4_from_form(h.f, chan1); return; }
  if (kind == 6 || kind == 7) {
    const HintFields hf = parse_hints(h.mode, h.hints);
    d.lo = eac_r11_from_form(h, hf, chan0, high_quality);
    if (kind == 7) d.hi = eac_r11_from_form(h, hf, chan1, high_quality);
    return;
  }
  const HintFields hf = parse_hints(h.mode, h.hints);
  if (kind == 2 && !high_quality && hf.bc1_hint0) {
    // hint0: the colour block comes from the endpoints and weights as they are -- no texels needed (BC1 1.80 -> 1.65 ms per
    // 2^24 blocks of C5).  (The same shortcut for BC3, with the alpha channel alone interpolated for its BC4 half, measured
    // slower: 2.09 -> 2.96 ms; BC3 keeps the full interpolation.)
    d.lo = bc1_bytes(bc1_from_hint0(h));
    return;
  }
  uint32_t px[16];
  texels_from_form(h.f, px);
  switch (kind) {
    case 0: d.lo = etc1_from_form(h, hf, px); break;
    case 1: d.lo = eac_a8_from_form(h, hf, px); d.hi = etc1_from_form(h, hf, px); break;
    case 2: d.lo = bc1_from_form(h, hf, px, high_quality); break;
    case 3: d.lo = bc4_encode(px, 3); d.hi = bc1_from_form(h, hf, px, high_quality); break;
    default: break;
  }
}

// One block, mode not known in advance (host checker, simple kernels).  Returns false for an invalid block (d zeroed).
UB_FN bool transcode_block_hinted(const Bits128& src, uint32_t kind, bool high_quality, uint32_t chan0, uint32_t chan1, Bits128& d) {
  HintedForm h;
  if (!hinted_form_t<-1>(src, h)) { d.lo = 0; d.hi = 0; return false; }
  hinted_target_from_form(h, kind, high_quality, chan0, chan1, d);
  return true;
}

}  // namespace ub
