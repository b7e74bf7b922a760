// uastc_modes.cuh -- per-mode candidate generation and partition estimation for the UASTC encoder.
// Restates astc_mode0_or_18 ... astc_mode15 (reference external/basis_universal/encoder/basisu_uastc_enc.cpp:471-2509),
// estimate_partition2 / _list (:639, :1543), the inline 3-subset and mode-7 estimators (:829-858, :1363-1405) and
// color_cell_compression_est_astc (encoder/basisu_bc7enc.cpp:1764-1984).
//
// The 14 reference functions collapse into two shapes: "N subsets, one plane" and "one subset, two planes"; each
// is a list of colour-cell jobs (texel subset, weight count, endpoint range, alpha) fed to color_cell_compression.
#pragma once
#include "uastc_ccc.cuh"
#include "uastc_decode.cuh"

namespace ub {

struct EncCfg {
  uint32_t level, flags, mode_mask, uber_level, ls_passes;
  bool estimate_partition, always_try_alpha, bc1_hints, only_la_on_transparent;
  bool cta_sync;  // device only: the whole CTA calls the encoder phases together and meets at barriers between stages
  uint32_t eac_mul_rad, eac_table_mask;
};

// level -> search effort (basisu_uastc_enc.cpp:3188-3256)
UB_FN EncCfg enc_cfg_from_flags(uint32_t flags) {
  EncCfg c;
  int level = (int)(flags & 7u);
  if (level > 4) level = 4;
  c.level = (uint32_t)level; c.flags = flags; c.cta_sync = false;
  c.mode_mask = 0xFFFFFFFFu; c.uber_level = 6; c.estimate_partition = false; c.always_try_alpha = true;
  c.eac_mul_rad = 3; c.eac_table_mask = 0xFFFFFFFFu; c.ls_passes = 2; c.bc1_hints = true; c.only_la_on_transparent = false;
  const uint32_t fast_tables = (1u << 2) | (1u << 8) | (1u << 11) | (1u << 13);
  switch (level) {
    case 0:
      c.mode_mask = (1u << 0) | (1u << 8) | (1u << 11) | (1u << 12) | (1u << 15);
      c.always_try_alpha = false; c.eac_mul_rad = 0; c.eac_table_mask = fast_tables; c.uber_level = 0; c.ls_passes = 1;
      c.bc1_hints = false; c.estimate_partition = true; c.only_la_on_transparent = true;
      break;
    case 1:
      c.mode_mask = (1u << 0) | (1u << 4) | (1u << 6) | (1u << 8) | (1u << 9) | (1u << 11) | (1u << 12) | (1u << 15) | (1u << 17);
      c.always_try_alpha = false; c.eac_mul_rad = 0; c.eac_table_mask = fast_tables; c.uber_level = 0; c.ls_passes = 1;
      c.estimate_partition = true;
      break;
    case 2:
      c.mode_mask = (1u << 0) | (1u << 1) | (1u << 4) | (1u << 5) | (1u << 6) | (1u << 8) | (1u << 9) | (1u << 10) | (1u << 11) |
                    (1u << 12) | (1u << 13) | (1u << 15) | (1u << 16) | (1u << 17);
      c.always_try_alpha = false; c.eac_mul_rad = 1;
      c.eac_table_mask = (1u << 0) | (1u << 2) | (1u << 6) | (1u << 7) | (1u << 8) | (1u << 10) | (1u << 11) | (1u << 13);
      c.uber_level = 1; c.ls_passes = 1; c.estimate_partition = true;
      break;
    case 3:
      c.always_try_alpha = false; c.eac_mul_rad = 2; c.uber_level = 3; c.estimate_partition = true;
      break;
    default: break;
  }
  return c;
}

// ---- fast partition error estimate (color_cell_compression_est_astc per subset, channel weights all 1): the subset's
// bounding box gives the axis and the two end colours, the palette is interpolated between them, every texel takes the
// palette entry whose projection on the axis is nearest, squared distances are summed.
UB_FN uint32_t minu4(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __vminu4(a, b);
#else
  uint32_t r = 0;
  for (uint32_t c = 0; c < 4; c++) { const uint32_t x = ch(a, c), y = ch(b, c); r |= (x < y ? x : y) << (8 * c); }
  return r;
#endif
}
UB_FN uint32_t maxu4(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __vmaxu4(a, b);
#else
  uint32_t r = 0;
  for (uint32_t c = 0; c < 4; c++) { const uint32_t x = ch(a, c), y = ch(b, c); r |= (x > y ? x : y) << (8 * c); }
  return r;
#endif
}

// total estimate of one pattern: `pat` = subset per texel (2 bits each).  All threads of a warp walk the pattern list
// together, so the membership tests below are uniform branches and the subsets need no gathering.
UB_FN uint64_t est_pattern(uint32_t nw, uint32_t ncomps, const uint32_t* px, uint32_t pat, uint32_t subsets) {
  const uint32_t cmask = (ncomps == 4) ? 0xFFFFFFFFu : 0x00FFFFFFu;   // RGB cells: alpha takes no part
  uint32_t total = 0;
#pragma unroll 1
  for (uint32_t s = 0; s < subsets; s++) {
    uint32_t lo = 0xFFFFFFFFu, hi = 0;
#pragma unroll
    for (uint32_t i = 0; i < 16; i++)
      if (((pat >> (2 * i)) & 3u) == s) { lo = minu4(lo, px[i]); hi = maxu4(hi, px[i]); }
    lo &= cmask; hi &= cmask;
    const uint32_t axis = hi - lo;   // per byte hi >= lo: no borrows
    // palette entries (astc_interpolate_linear per channel: (l*257*(64-w) + h*257*w + 32) >> 14) and, between
    // neighbours, the projection value from which the upper one is nearer
    uint32_t wc[8];
    int thr[8];
    int base[4], delta[4];
#pragma unroll
    for (uint32_t c = 0; c < 4; c++) {
      const int l = (int)(ch(lo, c) * 257u), h = (int)(ch(hi, c) * 257u);
      base[c] = l * 64 + 32; delta[c] = h - l;
    }
    int prev = 0;
#pragma unroll
    for (uint32_t i = 0; i < 8; i++) {
      if (i < nw) {
        uint32_t v;
        if (i == 0) v = lo;
        else if (i == nw - 1) v = hi;
        else {
          const int w = (int)sel_weight(nw, i);
          v = pk((uint32_t)((base[0] + delta[0] * w) >> 14), (uint32_t)((base[1] + delta[1] * w) >> 14),
                 (uint32_t)((base[2] + delta[2] * w) >> 14), (uint32_t)((base[3] + delta[3] * w) >> 14));
        }
        wc[i] = v;
        const int dot = (int)dot4u(v, axis);
        thr[i] = (prev + dot + 1) >> 1;   // thr[i]: boundary between entries i-1 and i (thr[0] unused)
        prev = dot;
      }
    }
#pragma unroll
    for (uint32_t i = 0; i < 16; i++)
      if (((pat >> (2 * i)) & 3u) == s) {
        const uint32_t c = px[i] & cmask;
        const int d = (int)dot4u(c, axis);
        // the projections of the palette ascend along the axis, so "the highest boundary not above d" is a count
        uint32_t sel = 0;
#pragma unroll
        for (uint32_t j = 1; j < 8; j++) if (j < nw && d >= thr[j]) sel++;
        total += sqdist4(wc[sel], c);
      }
  }
  return total;
}

// kind: 2 = the 30 two-subset patterns, 3 = the 11 three-subset patterns, 7 = the 19 mode-7 patterns
UB_FN uint32_t pattern_of(int kind, uint32_t idx) {
  if (kind == 3) return UBT(astc_pat3)[idx];
  return spread_mask2(kind == 7 ? UBT(astc_pat7)[idx] : UBT(astc_pat2)[idx]);
}
UB_FN uint32_t pattern_count(int kind) { return kind == 3 ? 11u : (kind == 7 ? 19u : 30u); }

// Best `max_parts` patterns, ascending error, first-wins on ties (estimate_partition2 == list of one).
UB_BIG void estimate_partitions(int kind, uint32_t nw, uint32_t ncomps, const uint32_t* px, uint32_t* parts, uint32_t max_parts) {
  uint64_t errs[8];
  for (uint32_t i = 0; i < max_parts; i++) { errs[i] = ~0ull; parts[i] = 0; }
  const uint32_t total = pattern_count(kind), subsets = (kind == 3) ? 3u : 2u;
  for (uint32_t cp = 0; cp < total; cp++) {
    const uint64_t e = est_pattern(nw, ncomps, px, pattern_of(kind, cp), subsets);
    for (uint32_t i = 0; i < max_parts; i++)
      if (e < errs[i]) {
        for (uint32_t j = max_parts - 1; j > i; --j) { parts[j] = parts[j - 1]; errs[j] = errs[j - 1]; }
        parts[i] = cp; errs[i] = e;
        break;
      }
  }
}

UB_FN uint32_t sum_unq_rgb(uint32_t range, const uint8_t* ep, uint32_t which) {
  return astc_unquant(range, ep[0 + which]) + astc_unquant(range, ep[2 + which]) + astc_unquant(range, ep[4 + which]);
}

// One colour cell's result as the candidate assembly needs it (24 bytes: what travels between the cell jobs and the
// candidate jobs of the job-based encode kernel).
struct CellOut {
  uint32_t astc_lo, astc_hi;   // endpoints as BISE codes, r | g<<8 | b<<16 | a<<24
  alignas(8) uint8_t sel[16];  // weights of the cell's texels, in texel order
};

UB_FN void candidate_init(uint32_t mode, uint32_t pattern, uint32_t ccs, Candidate& c) {
  c.mode = mode; c.pattern = pattern; c.ccs = ccs;
  for (uint32_t i = 0; i < 18; i++) c.ep[i] = 0;
  for (uint32_t i = 0; i < 32; i++) c.w[i] = 0;
}

// "N subsets, one plane": modes 0,1,2,3,4,5,7,9,10,12,14,15,16,18.  Subset s of the candidate = one cell job:
//   px     the 16 texels the cells see (already (l,0,0,a) for LA modes)
//   kind   partition family (0 = single subset)
UB_BIG void single_plane_cell(uint32_t mode, int kind, uint32_t pattern, uint32_t s, const uint32_t* px, const EncCfg& cfg, CellOut& o,
                             const uint8_t* force_sel = nullptr) {
  const uint32_t d = UBT(uastc_mode_desc)[mode];
  const uint32_t comps = md_comps(d), range = md_endpoint_range(d), nw = 1u << md_weight_bits(d);
  const uint32_t pat = kind ? pattern_of(kind, pattern) : 0u;
  CccParams p;
  p.nw = nw; p.range = range; p.has_alpha = (comps != 3); p.force_sel = force_sel;
  p.uber_level = cfg.uber_level; p.ls_passes = cfg.ls_passes;
  uint32_t sub[16], n = 0;
  for (uint32_t i = 0; i < 16; i++) if (((pat >> (2 * i)) & 3u) == s) { sub[n] = px[i]; n++; }
  p.n = n;
  CccResults r;
  color_cell_compression(p, sub, r);
  o.astc_lo = r.astc_lo; o.astc_hi = r.astc_hi;
#pragma unroll
  for (uint32_t i = 0; i < 4; i++) ((uint32_t*)o.sel)[i] = ((const uint32_t*)r.sel)[i];
}

// ... and its place in the candidate: endpoint codes of subset s (with the ASTC blue-contraction ordering: s1 < s0 swaps the
// endpoints and inverts the weights) and the weights of the subset's texels.
UB_FN void single_plane_place(uint32_t mode, int kind, uint32_t pattern, uint32_t s, const CellOut& r, Candidate& c, bool forced) {
  const uint32_t d = UBT(uastc_mode_desc)[mode];
  const uint32_t comps = md_comps(d), range = md_endpoint_range(d), nw = 1u << md_weight_bits(d);
  const uint32_t pat = kind ? pattern_of(kind, pattern) : 0u;
  uint8_t* ep = c.ep + s * comps * 2;
  if (comps == 2) {
    ep[0] = (uint8_t)ch(r.astc_lo, 0); ep[1] = (uint8_t)ch(r.astc_hi, 0);
    ep[2] = (uint8_t)ch(r.astc_lo, 3); ep[3] = (uint8_t)ch(r.astc_hi, 3);
  } else {
    for (uint32_t k = 0; k < comps; k++) { ep[2 * k] = (uint8_t)ch(r.astc_lo, k); ep[2 * k + 1] = (uint8_t)ch(r.astc_hi, k); }
  }
  bool invert = false;
  if (comps >= 3 && !forced) {
    if (sum_unq_rgb(range, ep, 1) < sum_unq_rgb(range, ep, 0)) {
      for (uint32_t k = 0; k < comps; k++) { const uint8_t t = ep[2 * k]; ep[2 * k] = ep[2 * k + 1]; ep[2 * k + 1] = t; }
      invert = true;
    }
  }
  uint32_t j = 0;
  for (uint32_t i = 0; i < 16; i++)
    if (((pat >> (2 * i)) & 3u) == s) { c.w[i] = (uint8_t)(invert ? (nw - 1 - r.sel[j]) : r.sel[j]); j++; }
}

UB_BIG void mode_single_plane(uint32_t mode, int kind, uint32_t pattern, const uint32_t* px, const EncCfg& cfg, Candidate& c,
                             const uint8_t* force_sel = nullptr) {
  const uint32_t subsets = md_subsets(UBT(uastc_mode_desc)[mode]);
  candidate_init(mode, pattern, 0, c);
  for (uint32_t s = 0; s < subsets; s++) {
    CellOut o;
    single_plane_cell(mode, kind, pattern, s, px, cfg, o, force_sel);
    single_plane_place(mode, kind, pattern, s, o, c, force_sel != nullptr);
  }
}

// "one subset, two planes": modes 6 (rot 0..2), 11 and 13 (rot 0..3), 17 (L plane + A plane).  Plane 0 = the colour
// plane with the rotated channel replaced, plane 1 = the rotated channel alone: two cell jobs of 16 texels.
UB_BIG void dual_plane_cell(uint32_t mode, uint32_t rot, uint32_t plane, const uint32_t* px, const EncCfg& cfg, CellOut& o) {
  const uint32_t d = UBT(uastc_mode_desc)[mode];
  const uint32_t range = md_endpoint_range(d), nw = 1u << md_weight_bits(d);
  uint32_t q[16];
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t v = px[i];
    if (mode == 17) {
      q[i] = ch(v, plane ? 3 : 0) * 0x010101u | 0xFF000000u;
    } else if (plane) {
      q[i] = ch(v, rot) * 0x010101u | 0xFF000000u;
    } else {
      uint32_t t = v;
      if (mode == 6) t = (t & ~(255u << (8 * rot))) | (255u << (8 * rot));
      else { if (rot < 3) t = (t & ~(255u << (8 * rot))) | (ch(v, 3) << (8 * rot)); t |= 0xFF000000u; }
      q[i] = t;
    }
  }
  CccParams p;
  p.n = 16; p.nw = nw; p.range = range; p.has_alpha = false; p.force_sel = nullptr;
  p.uber_level = cfg.uber_level; p.ls_passes = cfg.ls_passes;
  CccResults r;
  color_cell_compression(p, q, r);
  o.astc_lo = r.astc_lo; o.astc_hi = r.astc_hi;
#pragma unroll
  for (uint32_t i = 0; i < 4; i++) ((uint32_t*)o.sel)[i] = ((const uint32_t*)r.sel)[i];
}

UB_FN void dual_plane_assemble(uint32_t mode, uint32_t rot, const CellOut& r0, const CellOut& r1, Candidate& c) {
  const uint32_t d = UBT(uastc_mode_desc)[mode];
  const uint32_t comps = md_comps(d), range = md_endpoint_range(d), nw = 1u << md_weight_bits(d);
  c.mode = mode; c.pattern = 0; c.ccs = (mode == 17) ? 3u : rot;
  for (uint32_t i = 0; i < 18; i++) c.ep[i] = 0;
  bool invert = false;
  if (mode == 17) {
    c.ep[0] = (uint8_t)ch(r0.astc_lo, 0); c.ep[1] = (uint8_t)ch(r0.astc_hi, 0);
    c.ep[2] = (uint8_t)ch(r1.astc_lo, 0); c.ep[3] = (uint8_t)ch(r1.astc_hi, 0);
  } else {
    for (uint32_t k = 0; k < 3; k++) {
      const CellOut& rr = (rot == k) ? r1 : r0;
      c.ep[2 * k] = (uint8_t)ch(rr.astc_lo, k); c.ep[2 * k + 1] = (uint8_t)ch(rr.astc_hi, k);
    }
    if (comps == 4) {
      if (rot == 3) { c.ep[6] = (uint8_t)ch(r1.astc_lo, 0); c.ep[7] = (uint8_t)ch(r1.astc_hi, 0); }
      else { c.ep[6] = (uint8_t)ch(r0.astc_lo, rot); c.ep[7] = (uint8_t)ch(r0.astc_hi, rot); }
    }
    if (sum_unq_rgb(range, c.ep, 1) < sum_unq_rgb(range, c.ep, 0)) {
      for (uint32_t k = 0; k < comps; k++) { const uint8_t t = c.ep[2 * k]; c.ep[2 * k] = c.ep[2 * k + 1]; c.ep[2 * k + 1] = t; }
      invert = true;
    }
  }
  for (uint32_t i = 0; i < 16; i++) {
    c.w[2 * i] = (uint8_t)(invert ? (nw - 1 - r0.sel[i]) : r0.sel[i]);
    c.w[2 * i + 1] = (uint8_t)(invert ? (nw - 1 - r1.sel[i]) : r1.sel[i]);
  }
}

UB_BIG void mode_dual_plane(uint32_t mode, uint32_t rot, const uint32_t* px, const EncCfg& cfg, Candidate& c) {
  CellOut r0, r1;
  dual_plane_cell(mode, rot, 0, px, cfg, r0);
  dual_plane_cell(mode, rot, 1, px, cfg, r1);
  dual_plane_assemble(mode, rot, r0, r1, c);
}

}  // namespace ub
