// uastc_encode.cuh -- one 4x4 RGBA block -> one UASTC block.
// Restates encode_uastc (reference external/basis_universal/encoder/basisu_uastc_enc.cpp:3127-3646): classify,
// run the enabled modes, score every candidate by UASTC + transcoded-BC7 error, pick, compute hints, pack
// (pack_uastc :111-466).
#pragma once
#include "uastc_modes.cuh"
#include "uastc_hints.cuh"

namespace ub {

const uint32_t MAX_CANDIDATES = 512;

// Hints carried in the block besides the colour data.
struct BlockHints {
  uint32_t bc1_hint0, bc1_hint1;
  uint32_t etc1_flip, etc1_diff, etc1_inten0, etc1_inten1, etc1_bias;
  uint32_t etc2_hints;                               // eac table | multiplier << 4
  uint32_t etc1_selector, etc1_r, etc1_g, etc1_b;    // solid-colour blocks only
};

// Anchor normalisation (first half of pack_uastc, basisu_uastc_enc.cpp:262-338): every subset's (plane's) anchor
// weight must have a clear MSB; otherwise flip that subset's weights and swap its endpoints.
UB_BIG void normalize_candidate(const Candidate& c, Candidate& o) {
  o = c;
  if (c.mode == UASTC_MODE_SOLID) return;
  const uint32_t d = UBT(uastc_mode_desc)[c.mode];
  const uint32_t comps = md_comps(d), subsets = md_subsets(d), planes = 1 + md_dual_plane(d), wb = md_weight_bits(d);
  const uint32_t pat = cand_pat(c, d);
  uint32_t anchors = 0;
  if (subsets == 3) anchors = UBT(astc_anchor3)[c.pattern];
  else if (subsets == 2) anchors = (c.mode == 7) ? UBT(astc_anchor7)[c.pattern] : UBT(astc_anchor2)[c.pattern];
  const uint32_t wmax = (1u << wb) - 1u;
  for (uint32_t plane = 0; plane < planes; plane++)
    for (uint32_t s = 0; s < subsets; s++) {
      const uint32_t a = (anchors >> (4 * s)) & 15u;
      if (o.w[a * planes + plane] & (1u << (wb - 1))) {
        for (uint32_t i = 0; i < 16; i++)
          if (((pat >> (2 * i)) & 3u) == s) o.w[i * planes + plane] = (uint8_t)(wmax - o.w[i * planes + plane]);
        if (planes == 2) {
          for (uint32_t k = 0; k < comps; k++) {
            const uint32_t cp = (comps == 2) ? k : ((k == c.ccs) ? 1u : 0u);
            if (cp == plane) { const uint8_t t = o.ep[2 * k]; o.ep[2 * k] = o.ep[2 * k + 1]; o.ep[2 * k + 1] = t; }
          }
        } else {
          for (uint32_t k = 0; k < comps; k++) {
            const uint32_t i0 = s * comps * 2 + 2 * k;
            const uint8_t t = o.ep[i0]; o.ep[i0] = o.ep[i0 + 1]; o.ep[i0 + 1] = t;
          }
        }
      }
    }
}

// Bit emission (second half of pack_uastc, :111-260 and :340-466).  `c` must already be normalised.
UB_BIG Bits128 pack_uastc(const Candidate& c, uint32_t solid_rgba, const BlockHints& h) {
  const uint32_t mode = c.mode;
  const uint32_t d = UBT(uastc_mode_desc)[mode];
  BitWriter bw;
  bw.init();
  bw.put(md_huff_code(d), md_huff_len(d));
  if (mode == UASTC_MODE_SOLID) {
    bw.put(solid_rgba, 32);
    bw.put(h.etc1_diff, 1); bw.put(h.etc1_inten0, 3); bw.put(h.etc1_selector, 2);
    bw.put(h.etc1_r, 5); bw.put(h.etc1_g, 5); bw.put(h.etc1_b, 5);
    return bw.b;
  }
  if (UBT(uastc_mode_has_bc1_hint0)[mode]) bw.put(h.bc1_hint0, 1);
  if (UBT(uastc_mode_has_bc1_hint1)[mode]) bw.put(h.bc1_hint1, 1);
  bw.put(h.etc1_flip, 1); bw.put(h.etc1_diff, 1); bw.put(h.etc1_inten0, 3); bw.put(h.etc1_inten1, 3);
  if (UBT(uastc_mode_has_etc1_bias)[mode]) bw.put(h.etc1_bias, 5);
  const uint32_t comps = md_comps(d), subsets = md_subsets(d), planes = 1 + md_dual_plane(d), wb = md_weight_bits(d);
  if (comps != 3) bw.put(h.etc2_hints, 8);  // every mode with alpha or LA carries the ETC2 EAC hint byte
  if (subsets == 3) bw.put(c.pattern, 4); else if (subsets == 2) bw.put(c.pattern, 5);
  if (planes == 2 && mode != 17) bw.put(c.ccs, 2);
  const uint8_t* w = c.w;
  const uint8_t* ep = c.ep;
  uint32_t anchors = 0;
  if (subsets == 3) anchors = UBT(astc_anchor3)[c.pattern];
  else if (subsets == 2) anchors = (mode == 7) ? UBT(astc_anchor7)[c.pattern] : UBT(astc_anchor2)[c.pattern];

  // endpoints: trit / quint digits bundled first, then the plain bits
  const uint32_t total_values = comps * 2 * subsets, er = md_endpoint_range(d);
  const uint32_t eb = UBT(astc_bise_range)[er * 3], trits = UBT(astc_bise_range)[er * 3 + 1], quints = UBT(astc_bise_range)[er * 3 + 2];
  if (trits | quints) {
    const uint32_t base = trits ? 3u : 5u, full = trits ? 243u : 125u;
    uint32_t acc = 0, mul = 1;
    for (uint32_t i = 0; i < total_values; i++) {
      acc += (ep[i] >> eb) * mul;
      mul *= base;
      if (mul == full) { bw.put(acc, trits ? 8 : 7); acc = 0; mul = 1; }
    }
    if (mul > 1) {
      uint32_t nb;
      if (trits) nb = (mul == 3) ? 2 : (mul == 9) ? 4 : (mul == 27) ? 5 : 7;
      else nb = (mul == 5) ? 3 : 5;
      bw.put(acc, nb);
    }
  }
  for (uint32_t i = 0; i < total_values; i++) bw.put(ep[i] & ((1u << eb) - 1u), eb);

  for (uint32_t i = 0; i < 16 * planes; i++) {
    uint32_t nb = wb;
    const uint32_t texel = i / planes;
    for (uint32_t s = 0; s < subsets; s++)
      if (((anchors >> (4 * s)) & 15u) == texel) { nb--; break; }
    bw.put(w[i] & ((1u << nb) - 1u), nb);
  }
  return bw.b;
}

// All hints of a non-solid block (basisu_uastc_enc.cpp:3607-3637).  nc = normalised winning candidate.
UB_FN void compute_hints(const Candidate& nc, const uint32_t* px, const uint32_t* dec, const EncCfg& cfg, BlockHints& h, bool active = true) {
  if (cfg.bc1_hints) { if (active) compute_bc1_hints(nc, px, dec, h.bc1_hint0, h.bc1_hint1); UB_STAGE_SYNC(cfg); }
  if (active) {
    const uint32_t d = UBT(uastc_mode_desc)[nc.mode];
    h.etc2_hints = (md_comps(d) != 3) ? compute_eac_hint(dec, cfg.eac_mul_rad, cfg.eac_table_mask) : 0u;
  }
  UB_STAGE_SYNC(cfg);
  if (active) compute_etc1_hints(nc.mode, px, dec, cfg.level, cfg.flags, h.etc1_flip, h.etc1_diff, h.etc1_inten0, h.etc1_inten1, h.etc1_bias);
  UB_STAGE_SYNC(cfg);
}

UB_FN void block_error(const uint32_t* a, const uint32_t* b, uint64_t& rgb, uint64_t& rgba, uint64_t& la) {
  uint64_t er = 0, eg = 0, eb = 0, ea = 0;
  for (uint32_t i = 0; i < 16; i++) {
    const int dr = (int)ch(a[i], 0) - (int)ch(b[i], 0), dg = (int)ch(a[i], 1) - (int)ch(b[i], 1);
    const int db = (int)ch(a[i], 2) - (int)ch(b[i], 2), da = (int)ch(a[i], 3) - (int)ch(b[i], 3);
    er += (uint32_t)(dr * dr); eg += (uint32_t)(dg * dg); eb += (uint32_t)(db * db); ea += (uint32_t)(da * da);
  }
  la = er + ea; rgb = er + eg + eb; rgba = rgb + ea;
}

// candidate -> the transcoder's view of it (weights at full width; BC7 packing normalises its own anchors)
UB_FN void candidate_to_unpacked(const Candidate& c, UnpackedUastc& u) {
  u.mode = c.mode; u.desc = UBT(uastc_mode_desc)[c.mode]; u.pattern = c.pattern; u.ccs = c.ccs;
  u.pat = cand_pat(c, u.desc); u.anchors = 0; u.solid = 0; u.hints = 0;
  for (uint32_t i = 0; i < 18; i++) u.ep[i] = c.ep[i];
  const uint32_t wb = md_weight_bits(u.desc), n = 16 * (1 + md_dual_plane(u.desc));
  u.w.lo = 0; u.w.hi = 0;
  for (uint32_t i = 0; i < n; i++) bits_or(u.w, i * wb, c.w[i]);
}

UB_FN float u64_to_float(uint64_t v) {
#ifdef __CUDA_ARCH__
  return __ull2float_rn(v);
#else
  return (float)v;
#endif
}

// How many candidates a mode contributes (basisu_uastc_enc.cpp:3290-3351): one per tried partition / rotation.
UB_FN uint32_t mode_candidate_count(uint32_t mode, bool est) {
  switch (mode) {
    case 2: case 4: return est ? 1u : 30u;
    case 3: return est ? 1u : 11u;
    case 7: return est ? 1u : 19u;
    case 9: case 16: return est ? 4u : 30u;
    case 6: return 3u;
    case 11: case 13: return 4u;
    default: return 1u;
  }
}

// Candidate generation (basisu_uastc_enc.cpp:3290-3351).  Returns the count.
// The candidate LIST is in the reference's order -- LA modes 15..17, opaque modes 0..7 and 18, alpha modes 9..14 -- because
// selection keeps the first of equally good candidates.  The WORK is ordered differently: each mode knows its slot in
// the list up front, and modes of the three block classes that run the same routine share a stage (one-subset,
// partition estimate, partitioned, dual-plane), so that in a lock-step CTA (stage barriers, see UB_STAGE_SYNC) warps
// holding opaque blocks and warps holding alpha blocks work side by side instead of waiting out each other's stages.
// `active` = this thread has a (non-solid) block; every thread of a lock-step CTA walks the same stage list.
UB_BIG uint32_t generate_candidates(const uint32_t* px, bool has_alpha, bool is_la, const EncCfg& cfg, Candidate* cands, bool active) {
  const uint32_t mm = cfg.mode_mask;
  const bool est = cfg.estimate_partition;
  const bool la = active && is_la, opaque = active && !has_alpha, alpha = active && (has_alpha || cfg.always_try_alpha);
  uint32_t n = 0;
#define UB_SLOT(var, mode, group) const bool on##var = ((mm >> (mode)) & 1u) && (group); const uint32_t var = n; if (on##var) n += mode_candidate_count(mode, est);
  UB_SLOT(a15, 15, la) UB_SLOT(a16, 16, la) UB_SLOT(a17, 17, la)
  UB_SLOT(a0, 0, opaque) UB_SLOT(a1, 1, opaque) UB_SLOT(a2, 2, opaque) UB_SLOT(a3, 3, opaque) UB_SLOT(a4, 4, opaque)
  UB_SLOT(a5, 5, opaque) UB_SLOT(a6, 6, opaque) UB_SLOT(a7, 7, opaque) UB_SLOT(a18, 18, opaque)
  UB_SLOT(a9, 9, alpha) UB_SLOT(a10, 10, alpha) UB_SLOT(a11, 11, alpha) UB_SLOT(a12, 12, alpha) UB_SLOT(a13, 13, alpha) UB_SLOT(a14, 14, alpha)
#undef UB_SLOT
  uint32_t la_px[16];
  if (la) for (uint32_t i = 0; i < 16; i++) la_px[i] = ch(px[i], 0) | (px[i] & 0xFF000000u);

  // one subset, one plane
  if (ona15) mode_single_plane(15, 0, 0, la_px, cfg, cands[a15]);
  if (ona0) mode_single_plane(0, 0, 0, px, cfg, cands[a0]);
  if (ona10) mode_single_plane(10, 0, 0, px, cfg, cands[a10]);
  UB_STAGE_SYNC(cfg);
  if (ona1) mode_single_plane(1, 0, 0, px, cfg, cands[a1]);
  if (ona12) mode_single_plane(12, 0, 0, px, cfg, cands[a12]);
  UB_STAGE_SYNC(cfg);
  if (ona5) mode_single_plane(5, 0, 0, px, cfg, cands[a5]);
  if (ona14) mode_single_plane(14, 0, 0, px, cfg, cands[a14]);
  UB_STAGE_SYNC(cfg);
  // partition estimates
  uint32_t p16[4], p2[1], p3[1], p4[1], p7[1], p9[4];
  if (est) {
    if (ona16) estimate_partitions(2, 4, 4, la_px, p16, 4);
    if (ona2) estimate_partitions(2, 8, 3, px, p2, 1);
    if (ona9) estimate_partitions(2, 4, 4, px, p9, 4);
    UB_STAGE_SYNC(cfg);
    if (ona3) estimate_partitions(3, 4, 3, px, p3, 1);
    if (ona4) estimate_partitions(2, 4, 3, px, p4, 1);
    if (ona7) estimate_partitions(7, 4, 3, px, p7, 1);
    UB_STAGE_SYNC(cfg);
  }
  // several subsets, one plane
  if (ona16) { if (est) for (uint32_t k = 0; k < 4; k++) mode_single_plane(16, 2, p16[k], la_px, cfg, cands[a16 + k]);
               else for (uint32_t cp = 0; cp < 30; cp++) mode_single_plane(16, 2, cp, la_px, cfg, cands[a16 + cp]); }
  if (ona2) { if (est) mode_single_plane(2, 2, p2[0], px, cfg, cands[a2]);
              else for (uint32_t cp = 0; cp < 30; cp++) mode_single_plane(2, 2, cp, px, cfg, cands[a2 + cp]); }
  if (ona3) { if (est) mode_single_plane(3, 3, p3[0], px, cfg, cands[a3]);
              else for (uint32_t cp = 0; cp < 11; cp++) mode_single_plane(3, 3, cp, px, cfg, cands[a3 + cp]); }
  if (ona9) { if (est) for (uint32_t k = 0; k < 2; k++) mode_single_plane(9, 2, p9[k], px, cfg, cands[a9 + k]);
              else for (uint32_t cp = 0; cp < 15; cp++) mode_single_plane(9, 2, cp, px, cfg, cands[a9 + cp]); }
  UB_STAGE_SYNC(cfg);
  if (ona4) { if (est) mode_single_plane(4, 2, p4[0], px, cfg, cands[a4]);
              else for (uint32_t cp = 0; cp < 30; cp++) mode_single_plane(4, 2, cp, px, cfg, cands[a4 + cp]); }
  if (ona7) { if (est) mode_single_plane(7, 7, p7[0], px, cfg, cands[a7]);
              else for (uint32_t cp = 0; cp < 19; cp++) mode_single_plane(7, 7, cp, px, cfg, cands[a7 + cp]); }
  if (ona9) { if (est) for (uint32_t k = 2; k < 4; k++) mode_single_plane(9, 2, p9[k], px, cfg, cands[a9 + k]);
              else for (uint32_t cp = 15; cp < 30; cp++) mode_single_plane(9, 2, cp, px, cfg, cands[a9 + cp]); }
  UB_STAGE_SYNC(cfg);
  // two planes (the opaque side's last one-subset mode rides along with the alpha side's second dual-plane mode)
  if (ona17) mode_dual_plane(17, 0, px, cfg, cands[a17]);
  if (ona6) for (uint32_t rot = 0; rot < 3; rot++) mode_dual_plane(6, rot, px, cfg, cands[a6 + rot]);
  if (ona11) for (uint32_t rot = 0; rot < 4; rot++) mode_dual_plane(11, rot, px, cfg, cands[a11 + rot]);
  UB_STAGE_SYNC(cfg);
  if (ona18) mode_single_plane(18, 0, 0, px, cfg, cands[a18]);
  if (ona13) for (uint32_t rot = 0; rot < 4; rot++) mode_dual_plane(13, rot, px, cfg, cands[a13 + rot]);
  UB_STAGE_SYNC(cfg);
  return n;
}

// UASTC and transcoded-BC7 squared error of one candidate against the source texels (basisu_uastc_enc.cpp:3409-3456).
UB_BIG void score_candidate(const uint32_t* px, const Candidate& c, bool has_alpha, bool is_la, uint64_t& ue, uint64_t& be) {
  uint32_t dec[16];
  decode_uastc_pixels(c, dec);
  uint64_t u_rgb, u_rgba, u_la, b_rgb, b_rgba, b_la;
  block_error(px, dec, u_rgb, u_rgba, u_la);
  UnpackedUastc u;
  candidate_to_unpacked(c, u);
  Bc7Build bb;
  uastc_to_bc7_build(u, bb);
  decode_bc7_build(bb, dec);  // == decode_bc7_pixels(bc7_pack(bb)), see uastc_decode.cuh
  block_error(px, dec, b_rgb, b_rgba, b_la);
  ue = is_la ? u_la : (has_alpha ? u_rgba : u_rgb);
  be = is_la ? b_la : (has_alpha ? b_rgba : b_rgb);
}

// Candidate scoring + selection (basisu_uastc_enc.cpp:3389-3549).  Returns the winning index.
UB_BIG uint32_t select_candidate(const uint32_t* px, const Candidate* cands, uint32_t n, bool has_alpha, bool is_la, uint32_t flags,
                                uint64_t* overall, float* uerr_f) {
  if (n == 1) return 0;
  const bool favor_uastc = (flags & 8u) != 0;
  const bool favor_bc7 = !favor_uastc && ((flags & 16u) != 0);
  const uint32_t bc7_w = favor_bc7 ? 100u : (favor_uastc ? 0u : 50u);
  const uint32_t uastc_w = favor_bc7 ? 0u : 100u;
  double best_uerr = (double)1e+20f;
#pragma unroll 1
  for (uint32_t i = 0; i < n; i++) {
    uint64_t ue, be;
    score_candidate(px, cands[i], has_alpha, is_la, ue, be);
    overall[i] = (be * bc7_w) / 100 + (ue * uastc_w) / 100;
    if (!overall[i]) return i;
    uerr_f[i] = f_sqrt(u64_to_float(ue));
    if ((double)uerr_f[i] < best_uerr) best_uerr = (double)uerr_f[i];
  }
  const bool favor_simpler = (flags & 512u) != 0;  // cPackUASTCFavorSimplerModes (basisu_uastc_enc.h:62)
  uint64_t best_err = ~0ull;
  uint32_t best = 0;
  const bool all = (best_uerr == 0.0) || favor_bc7;
#pragma unroll 1
  for (uint32_t i = 0; i < n; i++) {
    if (!all) {
      const double delta = d_div((double)uerr_f[i], best_uerr);
      if (!(delta <= (double)1.3f)) continue;
    }
    const float wgt = (favor_simpler && (cands[i].mode == 0 || cands[i].mode == 10)) ? .8f : 1.0f;
    const uint64_t w = (uint64_t)f_mul(u64_to_float(overall[i]), wgt);
    if (w < best_err) { best_err = w; best = i; if (!best_err) break; }
  }
  return best;
}

// encode_uastc in three phases (candidates -> selection -> hints + pack).  `live` = this thread has a block; with
// cta_sync every thread of the CTA calls every phase (live or not) so the stage barriers inside line up.
struct EncState {
  bool solid, has_alpha, is_la;
  uint32_t n, best;
};

UB_FN void enc_phase_candidates(const uint32_t* px, uint32_t flags, Candidate* cands, EncState& st, bool live, bool cta_sync) {
  bool solid = true, has_alpha = false, is_la = true;
  if (live)
    for (uint32_t i = 0; i < 16; i++) {
      const uint32_t v = px[i];
      if ((v >> 24) < 255) has_alpha = true;
      if (v != px[0]) solid = false;
      if (ch(v, 0) != ch(v, 1) || ch(v, 0) != ch(v, 2)) is_la = false;
    }
  EncCfg cfg = enc_cfg_from_flags(flags);
  cfg.cta_sync = cta_sync;
  if (cfg.only_la_on_transparent && is_la && !has_alpha) is_la = false;
  st.solid = solid; st.has_alpha = has_alpha; st.is_la = is_la; st.best = 0;
  st.n = generate_candidates(px, has_alpha, is_la, cfg, cands, live && !solid);
}

UB_FN void enc_phase_select(const uint32_t* px, uint32_t flags, const Candidate* cands, uint64_t* overall, float* uerr_f, EncState& st, bool live) {
  if (live && !st.solid) st.best = select_candidate(px, cands, st.n, st.has_alpha, st.is_la, flags, overall, uerr_f);
}

UB_FN Bits128 enc_phase_finish(const uint32_t* px, uint32_t flags, const Candidate* cands, const EncState& st, bool live, bool cta_sync) {
  BlockHints h;
  memset(&h, 0, sizeof(h));
  EncCfg cfg = enc_cfg_from_flags(flags);
  cfg.cta_sync = cta_sync;
  const bool active = live && !st.solid;
  uint32_t dec[16];
  Candidate nc;
  if (active) {
    const Candidate& bc = cands[st.best];
    decode_uastc_pixels(bc, dec);
    normalize_candidate(bc, nc);
  }
  UB_STAGE_SYNC(cfg);
  compute_hints(nc, px, dec, cfg, h, active);
  Bits128 r;
  r.lo = 0; r.hi = 0;
  if (active) r = pack_uastc(nc, 0, h);
  else if (live) {
    Candidate c;
    memset(&c, 0, sizeof(c));
    c.mode = UASTC_MODE_SOLID;
    hints_solid_etc1(px[0], h.etc1_diff, h.etc1_inten0, h.etc1_selector, h.etc1_r, h.etc1_g, h.etc1_b);
    r = pack_uastc(c, px[0], h);
  }
  return r;
}

// encode_uastc.  px: 16 texels r | g<<8 | b<<16 | a<<24, raster order.
UB_FN Bits128 encode_uastc_block(const uint32_t* px, uint32_t flags, Candidate* cands, uint64_t* overall, float* uerr_f) {
  EncState st;
  enc_phase_candidates(px, flags, cands, st, true, false);
  enc_phase_select(px, flags, cands, overall, uerr_f, st, true);
  return enc_phase_finish(px, flags, cands, st, true, false);
}

}  // namespace ub
