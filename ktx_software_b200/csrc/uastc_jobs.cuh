// uastc_jobs.cuh -- encode_uastc at levels 0-3 (partition estimates on) cut into independent jobs.
//
// The reference runs one block at a time through a fixed list of modes (ENC/basisu_uastc_enc.cpp:3290-3351), each mode a
// handful of colour-cell fits, then scores every candidate and picks one (:3356-3549).  Here the same work is a set of
// job types whose instances are independent of each other, so that a CUDA CTA can hand them to its warps in any order
// and in homogeneous groups (uastc_b200.cu k_encode_jobs), while the host checker walks them one after the other
// (encode_uastc_block_jobs below, oracle/port_host.cpp):
//
//   estimate job   (block, estimate slot)   -> the partition patterns a partitioned mode will try      [estimate_partitions]
//   cell job       (block, cell slot)       -> endpoints + weights of one colour cell of one candidate [color_cell_compression]
//   candidate job  (block, candidate slot)  -> assemble the candidate from its cells, score it         [score_candidate]
//   block job      (block)                  -> select (first-wins rules of the reference), hints, pack
//
// Slots are static: candidate slot = position in the reference's candidate list if every mode were enabled and the
// block were of every class; a block's own list is the enabled slots of its class groups in ascending slot order,
// which is the reference's order.  Level 4 (no estimates, 170 candidates) has its own, arithmetic slot enumeration (j4_*, below).
#pragma once
#include "uastc_encode.cuh"

namespace ub {

const uint32_t J_GROUP_LA = 0, J_GROUP_OPQ = 1, J_GROUP_ALPHA = 2;
const uint32_t J_NUM_CAND = 32, J_NUM_CELL = 57, J_NUM_EST = 6;
const uint32_t J_LA_CANDS = 0x0000003Fu, J_OPQ_CANDS = 0x0001FFC0u, J_ALPHA_CANDS = 0xFFFE0000u;
const uint32_t J_MAX_CELL_SLOTS = 38, J_MAX_CAND_SLOTS = 21;   // per block: LA group + one of OPQ / ALPHA

// candidate slot -> mode | k << 5 | group << 8 | cells << 10 | first cell << 12
UB_TABLE(uint32_t, j_cand, 32) = {
#define JC(mode, k, g, n, c0) ((mode) | ((k) << 5) | ((g) << 8) | ((n) << 10) | ((c0) << 12))
  JC(15, 0, 0, 1, 0), JC(16, 0, 0, 2, 1), JC(16, 1, 0, 2, 3), JC(16, 2, 0, 2, 5), JC(16, 3, 0, 2, 7), JC(17, 0, 0, 2, 9),
  JC(0, 0, 1, 1, 11), JC(1, 0, 1, 1, 12), JC(2, 0, 1, 2, 13), JC(3, 0, 1, 3, 15), JC(4, 0, 1, 2, 18), JC(5, 0, 1, 1, 20),
  JC(6, 0, 1, 2, 21), JC(6, 1, 1, 2, 23), JC(6, 2, 1, 2, 25), JC(7, 0, 1, 2, 27), JC(18, 0, 1, 1, 29),
  JC(9, 0, 2, 2, 30), JC(9, 1, 2, 2, 32), JC(9, 2, 2, 2, 34), JC(9, 3, 2, 2, 36), JC(10, 0, 2, 1, 38),
  JC(11, 0, 2, 2, 39), JC(11, 1, 2, 2, 41), JC(11, 2, 2, 2, 43), JC(11, 3, 2, 2, 45), JC(12, 0, 2, 1, 47),
  JC(13, 0, 2, 2, 48), JC(13, 1, 2, 2, 50), JC(13, 2, 2, 2, 52), JC(13, 3, 2, 2, 54), JC(14, 0, 2, 1, 56)
#undef JC
};
UB_FN uint32_t jc_mode(uint32_t d) { return d & 31u; }
UB_FN uint32_t jc_k(uint32_t d) { return (d >> 5) & 7u; }
UB_FN uint32_t jc_group(uint32_t d) { return (d >> 8) & 3u; }
UB_FN uint32_t jc_cells(uint32_t d) { return (d >> 10) & 3u; }
UB_FN uint32_t jc_cell0(uint32_t d) { return d >> 12; }

// cell slot -> candidate slot | cell index << 5
UB_TABLE(uint8_t, j_cell, 57) = {
  0, 1, 1 | 32, 2, 2 | 32, 3, 3 | 32, 4, 4 | 32, 5, 5 | 32,
  6, 7, 8, 8 | 32, 9, 9 | 32, 9 | 64, 10, 10 | 32, 11, 12, 12 | 32, 13, 13 | 32, 14, 14 | 32, 15, 15 | 32, 16,
  17, 17 | 32, 18, 18 | 32, 19, 19 | 32, 20, 20 | 32, 21, 22, 22 | 32, 23, 23 | 32, 24, 24 | 32, 25, 25 | 32, 26,
  27, 27 | 32, 28, 28 | 32, 29, 29 | 32, 30, 30 | 32, 31};

// estimate slot -> mode | kind << 5 | weights << 8 | comps << 12 | patterns kept << 15 | first entry in the block's pattern list << 18 | group << 22
UB_TABLE(uint32_t, j_est, 6) = {
#define JE(mode, kind, nw, comps, keep, p0, g) ((mode) | ((kind) << 5) | ((nw) << 8) | ((comps) << 12) | ((keep) << 15) | ((p0) << 18) | ((g) << 22))
  JE(16, 2, 4, 4, 4, 0, 0), JE(2, 2, 8, 3, 1, 4, 1), JE(3, 3, 4, 3, 1, 5, 1), JE(4, 2, 4, 3, 1, 6, 1), JE(7, 7, 4, 3, 1, 7, 1), JE(9, 2, 4, 4, 4, 8, 2)
#undef JE
};
const uint32_t J_PARTS = 12;   // pattern list of a block: mode 16 x4, 2, 3, 4, 7, mode 9 x4

// scratch slots of a block: the LA group's slots first, then the block's one other group
UB_FN uint32_t j_cell_local(uint32_t gc) { return gc < 30u ? gc : gc - 19u; }
UB_FN uint32_t j_cand_local(uint32_t gs) { return gs < 17u ? gs : gs - 11u; }

// candidate slots whose mode the level's mode mask enables
UB_FN uint32_t j_enabled_cands(uint32_t mode_mask) {
  uint32_t m = 0;
  for (uint32_t gs = 0; gs < J_NUM_CAND; gs++) if ((mode_mask >> jc_mode(UBT(j_cand)[gs])) & 1u) m |= 1u << gs;
  return m;
}

// Block class: 0 = no block, J_CLS_SOLID, else J_CLS_LIVE | J_CLS_LA? | J_CLS_ALPHA?
const uint32_t J_CLS_LIVE = 1, J_CLS_LA = 2, J_CLS_ALPHA = 4, J_CLS_SOLID = 8;
UB_FN uint32_t j_classify(const uint32_t* px, const EncCfg& cfg) {
  bool solid = true, has_alpha = false, is_la = true;
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t v = px[i];
    if ((v >> 24) < 255) has_alpha = true;
    if (v != px[0]) solid = false;
    if (ch(v, 0) != ch(v, 1) || ch(v, 0) != ch(v, 2)) is_la = false;
  }
  if (solid) return J_CLS_SOLID;
  if (cfg.only_la_on_transparent && is_la && !has_alpha) is_la = false;
  return J_CLS_LIVE | (is_la ? J_CLS_LA : 0u) | (has_alpha ? J_CLS_ALPHA : 0u);
}
// candidate slots of a block of class cls (levels 0-3: alpha modes only for blocks with alpha)
UB_FN uint32_t j_block_cands(uint32_t cls, uint32_t enabled) {
  return enabled & (((cls & J_CLS_LA) ? J_LA_CANDS : 0u) | ((cls & J_CLS_ALPHA) ? J_ALPHA_CANDS : J_OPQ_CANDS));
}

UB_FN void j_la_texels(const uint32_t* px, uint32_t* la_px) {
  for (uint32_t i = 0; i < 16; i++) la_px[i] = ch(px[i], 0) | (px[i] & 0xFF000000u);
}

// pattern index of candidate slot gs in the block's pattern list, or 0 for unpartitioned modes
UB_FN uint32_t j_cand_pattern(uint32_t gs, const uint8_t* parts) {
  const uint32_t d = UBT(j_cand)[gs], mode = jc_mode(d), k = jc_k(d);
  switch (mode) {
    case 16: return parts[k];
    case 2: return parts[4];
    case 3: return parts[5];
    case 4: return parts[6];
    case 7: return parts[7];
    case 9: return parts[8 + k];
    default: return 0;
  }
}
UB_FN int j_mode_kind(uint32_t mode) { return (mode == 3) ? 3 : (mode == 7) ? 7 : (mode == 2 || mode == 4 || mode == 9 || mode == 16) ? 2 : 0; }
UB_FN bool j_mode_dual(uint32_t mode) { return mode == 6 || mode == 11 || mode == 13 || mode == 17; }

// ---- estimate job
UB_FN void j_run_estimate(uint32_t e, const uint32_t* px, uint8_t* parts) {
  const uint32_t d = UBT(j_est)[e];
  const uint32_t mode = d & 31u, keep = (d >> 15) & 7u, p0 = (d >> 18) & 15u;
  uint32_t la_px[16], out[4];
  if (mode == 16) j_la_texels(px, la_px);
  estimate_partitions((int)((d >> 5) & 7u), (d >> 8) & 15u, (d >> 12) & 7u, mode == 16 ? la_px : px, out, keep);
  for (uint32_t i = 0; i < keep; i++) parts[p0 + i] = (uint8_t)out[i];
}

// ---- cell job
UB_FN void j_run_cell(uint32_t gc, const uint32_t* px, const uint8_t* parts, const EncCfg& cfg, CellOut& o) {
  const uint32_t gs = UBT(j_cell)[gc] & 31u, cell = UBT(j_cell)[gc] >> 5;
  const uint32_t d = UBT(j_cand)[gs], mode = jc_mode(d);
  if (j_mode_dual(mode)) { dual_plane_cell(mode, jc_k(d), cell, px, cfg, o); return; }
  uint32_t la_px[16];
  const bool la = (mode == 15 || mode == 16);
  if (la) j_la_texels(px, la_px);
  single_plane_cell(mode, j_mode_kind(mode), j_cand_pattern(gs, parts), cell, la ? la_px : px, cfg, o);
}

// ---- candidate job: cells[0 .. jc_cells) of candidate slot gs -> the candidate
UB_FN void j_assemble(uint32_t gs, const uint8_t* parts, const CellOut* cells, Candidate& c) {
  const uint32_t d = UBT(j_cand)[gs], mode = jc_mode(d);
  if (j_mode_dual(mode)) { dual_plane_assemble(mode, jc_k(d), cells[0], cells[1], c); return; }
  const int kind = j_mode_kind(mode);
  const uint32_t pattern = j_cand_pattern(gs, parts), n = jc_cells(d);
  candidate_init(mode, pattern, 0, c);
  for (uint32_t s = 0; s < n; s++) single_plane_place(mode, kind, pattern, s, cells[s], c, false);
}

// score of a candidate as selection consumes it: overall error (fits 32 bits: <= 2 * 16 * 4 * 255^2) and sqrt(UASTC error)
UB_FN void j_score(const uint32_t* px, const Candidate& c, uint32_t cls, uint32_t flags, uint32_t& overall, float& uerr) {
  const bool favor_uastc = (flags & 8u) != 0;
  const bool favor_bc7 = !favor_uastc && ((flags & 16u) != 0);
  const uint32_t bc7_w = favor_bc7 ? 100u : (favor_uastc ? 0u : 50u);
  const uint32_t uastc_w = favor_bc7 ? 0u : 100u;
  uint64_t ue, be;
  score_candidate(px, c, (cls & J_CLS_ALPHA) != 0, (cls & J_CLS_LA) != 0, ue, be);
  overall = (uint32_t)((be * bc7_w) / 100 + (ue * uastc_w) / 100);
  uerr = f_sqrt(u64_to_float(ue));
}

// ---- level 4 ("full"): no partition estimates -- every pattern of every partitioned mode and every rotation of every dual-plane
// mode is a candidate (ENC:3290-3351 with estimate_partition off): 170 candidate slots and 343 cell slots, enumerated mode by
// mode in the reference's list order.  Alpha modes are tried on opaque blocks too (always_try_alpha, ENC:3201-3256), so the
// "alpha" group is every live block.  Within a warp-job all lanes share the mode AND the pattern: subsets have the same size in
// every lane, which the block-at-a-time level-4 kernel (evaluate_solution at 13.7 of 32 lanes) could not offer.
const uint32_t J4_NUM_MODES = 18, J4_NUM_CAND = 170, J4_NUM_CELL = 343;
// mode, candidates, cells per candidate, group
UB_TABLE(uint8_t, j4_modes, 72) = {
  15, 1, 1, 0,  16, 30, 2, 0,  17, 1, 2, 0,
  0, 1, 1, 1,  1, 1, 1, 1,  2, 30, 2, 1,  3, 11, 3, 1,  4, 30, 2, 1,  5, 1, 1, 1,  6, 3, 2, 1,  7, 19, 2, 1,  18, 1, 1, 1,
  9, 30, 2, 2,  10, 1, 1, 2,  11, 4, 2, 2,  12, 1, 1, 2,  13, 4, 2, 2,  14, 1, 1, 2};
UB_TABLE(uint16_t, j4_cand_base, 19) = {0, 1, 31, 32, 33, 34, 64, 75, 105, 106, 109, 128, 129, 159, 160, 164, 165, 169, 170};
UB_TABLE(uint16_t, j4_cell_base, 19) = {0, 1, 61, 63, 64, 65, 125, 158, 218, 219, 225, 263, 264, 324, 325, 333, 334, 342, 343};
UB_FN uint32_t j4_mode(uint32_t e) { return UBT(j4_modes)[e * 4]; }
UB_FN uint32_t j4_count(uint32_t e) { return UBT(j4_modes)[e * 4 + 1]; }
UB_FN uint32_t j4_cells(uint32_t e) { return UBT(j4_modes)[e * 4 + 2]; }
UB_FN uint32_t j4_group(uint32_t e) { return UBT(j4_modes)[e * 4 + 3]; }
UB_FN bool j4_block_in_group(uint32_t cls, uint32_t g) {
  return g == J_GROUP_LA ? (cls & J_CLS_LA) != 0 : (g == J_GROUP_OPQ ? !(cls & J_CLS_ALPHA) : true);
}
// cell `cell` of candidate k of mode entry e
UB_FN void j4_run_cell(uint32_t e, uint32_t k, uint32_t cell, const uint32_t* px, const EncCfg& cfg, CellOut& o) {
  const uint32_t mode = j4_mode(e);
  if (j_mode_dual(mode)) { dual_plane_cell(mode, k, cell, px, cfg, o); return; }
  uint32_t la_px[16];
  const bool la = (mode == 15 || mode == 16);
  if (la) j_la_texels(px, la_px);
  single_plane_cell(mode, j_mode_kind(mode), k, cell, la ? la_px : px, cfg, o);
}
UB_FN void j4_assemble(uint32_t e, uint32_t k, const CellOut* cells, Candidate& c) {
  const uint32_t mode = j4_mode(e);
  if (j_mode_dual(mode)) { dual_plane_assemble(mode, k, cells[0], cells[1], c); return; }
  const int kind = j_mode_kind(mode);
  const uint32_t n = j4_cells(e);
  candidate_init(mode, kind ? k : 0u, 0, c);
  for (uint32_t s = 0; s < n; s++) single_plane_place(mode, kind, kind ? k : 0u, s, cells[s], c, false);
}

// ---- block job, first half: the reference's selection rules (ENC:3456-3549) over the stored scores of the block's n
// candidates (list order); modes[i] = mode of candidate i.  Returns the winning list index.
UB_FN uint32_t j_select(uint32_t n, const uint32_t* overall, const float* uerr_f, const uint8_t* modes, uint32_t flags) {
  if (n == 1) return 0;
  const bool favor_uastc = (flags & 8u) != 0;
  const bool favor_bc7 = !favor_uastc && ((flags & 16u) != 0);
  double best_uerr = (double)1e+20f;
  for (uint32_t i = 0; i < n; i++) {
    if (!overall[i]) return i;   // the reference stops scoring at the first candidate with no error at all
    if ((double)uerr_f[i] < best_uerr) best_uerr = (double)uerr_f[i];
  }
  const bool favor_simpler = (flags & 512u) != 0;
  uint64_t best_err = ~0ull;
  uint32_t best = 0;
  const bool all = (best_uerr == 0.0) || favor_bc7;
  for (uint32_t i = 0; i < n; i++) {
    if (!all) {
      const double delta = d_div((double)uerr_f[i], best_uerr);
      if (!(delta <= (double)1.3f)) continue;
    }
    const float wgt = (favor_simpler && (modes[i] == 0 || modes[i] == 10)) ? .8f : 1.0f;
    const uint64_t w = (uint64_t)f_mul(u64_to_float((uint64_t)overall[i]), wgt);
    if (w < best_err) { best_err = w; best = i; if (!best_err) break; }
  }
  return best;
}

// ---- block job, second half: hints + pack of the winner (or of a solid block).  `bc` may be null when !active.
// cta_sync: every thread of the CTA calls this together (stage barriers inside).
UB_FN Bits128 j_finish(const uint32_t* px, uint32_t flags, const Candidate* bc, bool live, bool solid, bool cta_sync) {
  BlockHints h;
  memset(&h, 0, sizeof(h));
  EncCfg cfg = enc_cfg_from_flags(flags);
  cfg.cta_sync = cta_sync;
  const bool active = live && !solid;
  uint32_t dec[16];
  Candidate nc;
  if (active) {
    decode_uastc_pixels(*bc, dec);
    normalize_candidate(*bc, nc);
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

// ---- the block job split around the ETC1 hint jobs.  j_finish_head does everything of j_finish except the ETC1 hint search:
// the block comes out packed with the ETC1 hint fields (flip, diff, inten0, inten1[, bias]: one contiguous bit field right
// behind the mode code and the BC1 hint bits, ENC pack_uastc :111-135) left zero, `dec` holds the decoded winner and
// `etc1_at` the bit position of that field | has_bias << 8 (0xFFFF: nothing to fill in, solid or dead block).
UB_FN Bits128 j_finish_head(const uint32_t* px, uint32_t flags, const Candidate* bc, bool live, bool solid, bool cta_sync, uint32_t* dec,
                            uint32_t& mode, uint32_t& etc1_at) {
  BlockHints h;
  memset(&h, 0, sizeof(h));
  EncCfg cfg = enc_cfg_from_flags(flags);
  cfg.cta_sync = cta_sync;
  const bool active = live && !solid;
  Candidate nc;
  mode = UASTC_MODE_SOLID; etc1_at = 0xFFFFu;
  if (active) {
    decode_uastc_pixels(*bc, dec);
    normalize_candidate(*bc, nc);
    mode = nc.mode;
  }
  UB_STAGE_SYNC(cfg);
  if (cfg.bc1_hints) { if (active) compute_bc1_hints(nc, px, dec, h.bc1_hint0, h.bc1_hint1); UB_STAGE_SYNC(cfg); }
  Bits128 r;
  r.lo = 0; r.hi = 0;
  if (active) {
    const uint32_t d = UBT(uastc_mode_desc)[nc.mode];
    h.etc2_hints = (md_comps(d) != 3) ? compute_eac_hint(dec, cfg.eac_mul_rad, cfg.eac_table_mask) : 0u;
    r = pack_uastc(nc, 0, h);
    etc1_at = (md_huff_len(d) + UBT(uastc_mode_has_bc1_hint0)[nc.mode] + UBT(uastc_mode_has_bc1_hint1)[nc.mode]) |
              ((uint32_t)UBT(uastc_mode_has_etc1_bias)[nc.mode] << 8);
  } else if (live) {
    Candidate c;
    memset(&c, 0, sizeof(c));
    c.mode = UASTC_MODE_SOLID;
    hints_solid_etc1(px[0], h.etc1_diff, h.etc1_inten0, h.etc1_selector, h.etc1_r, h.etc1_g, h.etc1_b);
    r = pack_uastc(c, px[0], h);
  }
  return r;
}
// ... and the tail: the ETC1 hint jobs' results (in (flip slot, colour mode) order) into the packed block.
UB_FN void j_finish_tail(Bits128& r, uint32_t etc1_at, const Etc1JobOut* jobs, uint32_t n) {
  if (etc1_at == 0xFFFFu) return;
  uint32_t flip, diff, i0, i1, bias;
  etc1_hints_combine(jobs, n, flip, diff, i0, i1, bias);
  uint32_t f = flip | (diff << 1) | (i0 << 2) | (i1 << 5);
  if (etc1_at >> 8) f |= bias << 8;
  r.lo |= (uint64_t)f << (etc1_at & 255u);   // the field ends below bit 64 in every mode (7 + 2 + 13 bits at most)
}

// The jobs of one block run one after the other: what the CUDA kernel computes, in the order a single thread would.
// (Host checker; levels 0-3 only.)
UB_FN Bits128 j_etc1_jobs_and_tail(const uint32_t* px, const EncCfg& cfg, Bits128 r, const uint32_t* dec, uint32_t mode, uint32_t etc1_at) {
  const uint32_t js = etc1_job_slots(cfg.level, cfg.flags), nf = js & 3u, ni = js >> 2;
  Etc1JobOut jobs[4];
  for (uint32_t f = 0; f < nf; f++)
    for (uint32_t i = 0; i < ni; i++) etc1_hints_job(mode, px, dec, cfg.level, cfg.flags, f, i, jobs[f * ni + i]);
  j_finish_tail(r, etc1_at, jobs, nf * ni);
  return r;
}

// level 4: the jobs of k_encode_jobs<.., true>, one block after the other
UB_FN Bits128 encode_uastc_block_jobs_full(const uint32_t* px, uint32_t flags) {
  const EncCfg cfg = enc_cfg_from_flags(flags);
  const uint32_t cls = j_classify(px, cfg);
  if (cls == J_CLS_SOLID) return j_finish(px, flags, nullptr, true, true, false);
  uint32_t overall[J4_NUM_CAND], n = 0;
  float uerr[J4_NUM_CAND];
  uint8_t modes[J4_NUM_CAND], ent[J4_NUM_CAND], kk[J4_NUM_CAND];
  Candidate c;
  CellOut cells[3];
  for (uint32_t e = 0; e < J4_NUM_MODES; e++) {
    if (!j4_block_in_group(cls, j4_group(e))) continue;
    for (uint32_t k = 0; k < j4_count(e); k++) {
      for (uint32_t s = 0; s < j4_cells(e); s++) j4_run_cell(e, k, s, px, cfg, cells[s]);
      j4_assemble(e, k, cells, c);
      j_score(px, c, cls, flags, overall[n], uerr[n]);
      modes[n] = (uint8_t)j4_mode(e); ent[n] = (uint8_t)e; kk[n] = (uint8_t)k; n++;
    }
  }
  const uint32_t w = j_select(n, overall, uerr, modes, flags);
  for (uint32_t s = 0; s < j4_cells(ent[w]); s++) j4_run_cell(ent[w], kk[w], s, px, cfg, cells[s]);
  j4_assemble(ent[w], kk[w], cells, c);
  uint32_t dec[16], mode, etc1_at;
  const Bits128 r = j_finish_head(px, flags, &c, true, false, false, dec, mode, etc1_at);
  return j_etc1_jobs_and_tail(px, cfg, r, dec, mode, etc1_at);
}

UB_FN Bits128 encode_uastc_block_jobs(const uint32_t* px, uint32_t flags) {
  const EncCfg cfg = enc_cfg_from_flags(flags);
  if (cfg.level >= 4) return encode_uastc_block_jobs_full(px, flags);
  const uint32_t cls = j_classify(px, cfg);
  if (cls == J_CLS_SOLID) return j_finish(px, flags, nullptr, true, true, false);
  const uint32_t enabled = j_enabled_cands(cfg.mode_mask);
  const uint32_t mine = j_block_cands(cls, enabled);
  uint8_t parts[J_PARTS];
  for (uint32_t i = 0; i < J_PARTS; i++) parts[i] = 0;
  for (uint32_t e = 0; e < J_NUM_EST; e++) {
    const uint32_t d = UBT(j_est)[e], mode = d & 31u, g = d >> 22;
    const bool have = (g == J_GROUP_LA) ? (cls & J_CLS_LA) != 0 : ((g == J_GROUP_ALPHA) == ((cls & J_CLS_ALPHA) != 0));
    if (have && ((cfg.mode_mask >> mode) & 1u)) j_run_estimate(e, px, parts);
  }
  CellOut cells[J_NUM_CELL];
  for (uint32_t gc = 0; gc < J_NUM_CELL; gc++)
    if ((mine >> (UBT(j_cell)[gc] & 31u)) & 1u) j_run_cell(gc, px, parts, cfg, cells[gc]);
  uint32_t overall[J_MAX_CAND_SLOTS], slot[J_MAX_CAND_SLOTS], n = 0;
  float uerr[J_MAX_CAND_SLOTS];
  uint8_t modes[J_MAX_CAND_SLOTS];
  Candidate c;
  for (uint32_t gs = 0; gs < J_NUM_CAND; gs++)
    if ((mine >> gs) & 1u) {
      j_assemble(gs, parts, cells + jc_cell0(UBT(j_cand)[gs]), c);
      j_score(px, c, cls, flags, overall[n], uerr[n]);
      modes[n] = (uint8_t)jc_mode(UBT(j_cand)[gs]); slot[n] = gs; n++;
    }
  const uint32_t gs = slot[j_select(n, overall, uerr, modes, flags)];
  j_assemble(gs, parts, cells + jc_cell0(UBT(j_cand)[gs]), c);
  uint32_t dec[16], mode, etc1_at;
  const Bits128 r = j_finish_head(px, flags, &c, true, false, false, dec, mode, etc1_at);
  return j_etc1_jobs_and_tail(px, cfg, r, dec, mode, etc1_at);
}

}  // namespace ub
