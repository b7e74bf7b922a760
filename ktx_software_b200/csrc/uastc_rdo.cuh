// uastc_rdo.cuh -- UASTC rate-distortion post-pass, the per-block pieces.
// Restates uastc_rdo_blocks / uastc_recompute_hints (reference external/basis_universal/encoder/basisu_uastc_enc.cpp
// :3648-3728 recompute hints, :3730-3737 selector bit ranges, :3776-3792 match cost, :3825-4103 the block loop).
//
// The reference walks the blocks of a job range in order; for each block it splices the selector bits of up to
// dict/16 earlier blocks into the current block, scores every trial (UASTC error + transcoded BC7 error), and keeps
// the one minimising  ms_err * smooth_scale + lz_bits * lambda.  A map "selector bits -> latest block that carries
// them" prices matches and suppresses duplicate trials.  Only the trial loop is inherently parallel; the walk itself
// is a serial chain per job range.  The drivers (CUDA: uastc_b200.cu k_rdo_chain; host checker: oracle/port_host.cpp)
// share everything in this file.
#pragma once
#include "uastc_encode.cuh"

namespace ub {

struct RdoParams {
  float lambda;              // m_lambda
  float max_rms_ratio;       // m_max_allowed_rms_increase_ratio (10)
  float skip_rms;            // m_skip_block_rms_thresh (8)
  float smooth_scale;        // m_smooth_block_max_error_scale (10)
  float smooth_std_dev;      // m_max_smooth_block_std_dev (18)
  uint32_t window;           // max(1, m_lz_dict_size / 16) earlier blocks tried
  uint32_t literal_cost;     // m_lz_literal_cost (100)
  uint32_t flags;            // pack flags, for the hint recomputation
  uint32_t refine;           // m_endpoint_refinement
};

// g_uastc_mode_selector_bits (:3730): first selector bit and selector bit count of every mode
UB_TABLE(uint8_t, rdo_sel_bits, 38) = {65, 63, 69, 31, 73, 46, 89, 29, 89, 30, 68, 47, 66, 62, 89, 30, 0, 0, 97, 30,
                                       65, 63, 66, 62, 81, 47, 94, 30, 92, 31, 62, 63, 98, 30, 61, 62, 49, 79};

UB_FN uint32_t ilog2(uint32_t v) {  // v > 0
#ifdef __CUDA_ARCH__
  return 31u - (uint32_t)__clz((int)v);
#else
  return 31u - (uint32_t)__builtin_clz(v);
#endif
}

// compute_match_cost_estimate (:3776): 7 length bits + 5 distance bits + the deflate distance code's extra bits.
// The two reference tables are floor(log2) staircases: small[d] = log2(d)-1 (d >= 4), large[d>>8] = log2(d>>8)+7.
UB_FN uint32_t rdo_match_cost(uint32_t dist) {
  uint32_t extra;
  if (dist < 512) extra = (dist < 4) ? 0u : ilog2(dist) - 1u;
  else {
    const uint32_t d = dist < 32767u ? dist : 32767u;
    extra = ilog2(d >> 8) + 7u;
    while (dist >= 32768u) { extra++; dist >>= 1; }
  }
  return 12u + extra;
}

// bits [first, first+total) of a block, as (low <= 64 bits, remaining high bits)
UB_FN void rdo_get_sel(const Bits128& b, uint32_t first, uint32_t total, uint64_t& lo, uint64_t& hi) {
  lo = bits_get64(b, first);
  hi = 0;
  if (total < 64) lo &= lowmask64(total);
  else if (total > 64) hi = bits_get64(b, first + 64) & lowmask64(total - 64);
}

// the current block with the selector field replaced by another block's bits at the same position
UB_FN Bits128 rdo_splice(const Bits128& cur, const Bits128& prev, uint32_t first, uint32_t total) {
  Bits128 m;  // mask of the field
  m.lo = 0; m.hi = 0;
  if (first < 64) {
    m.lo = ~0ull << first;
    const uint32_t end = first + total;
    if (end < 64) m.lo &= lowmask64(end);
    else m.hi = lowmask64(end - 64);
  } else {
    m.hi = (~0ull << (first - 64)) & lowmask64(first + total - 64);
  }
  Bits128 r;
  r.lo = (cur.lo & ~m.lo) | (prev.lo & m.lo);
  r.hi = (cur.hi & ~m.hi) | (prev.hi & m.hi);
  return r;
}

UB_FN void unpacked_to_candidate(const UnpackedUastc& u, Candidate& c) {
  c.mode = u.mode; c.pattern = u.pattern; c.ccs = u.ccs;
  for (uint32_t i = 0; i < 18; i++) c.ep[i] = u.ep[i];
  const uint32_t n = 16u * (1u + md_dual_plane(u.desc));
  for (uint32_t i = 0; i < 32; i++) c.w[i] = (i < n) ? (uint8_t)uastc_weight(u, i) : (uint8_t)0;
}

// (UASTC error + transcoded-BC7 error) / 2 of a packed block against the source texels, RGBA squared distance
// (:3867-3893 and :3966-3990).  Returns false if the block does not unpack.
UB_BIG bool rdo_block_error(const uint32_t* px, const Bits128& blk, uint64_t& err, UnpackedUastc* out_u) {
  UnpackedUastc u;
  if (!unpack_uastc(blk, u, false)) return false;
  Candidate c;
  unpacked_to_candidate(u, c);
  uint64_t ue, be;
  score_candidate(px, c, true, false, ue, be);
  err = (ue + be) / 2;
  if (out_u) *out_u = u;
  return true;
}

// tracked_stat::get_std_dev over the 16 texels of one channel (basisu_math.h:2493)
UB_FN float rdo_channel_std_dev(const uint32_t* px, uint32_t c) {
  int64_t total = 0, total2 = 0;
  for (uint32_t i = 0; i < 16; i++) { const int64_t v = (int64_t)ch(px[i], c); total += v; total2 += v * v; }
  const int64_t var = 16 * total2 - total * total;
#ifdef __CUDA_ARCH__
  const float fv = __ll2float_rn(var);
#else
  const float fv = (float)var;
#endif
  return f_div(f_sqrt(fv), 16.0f);
}

// smooth_block_error_scale of a block (:3851-3858)
UB_FN float rdo_smooth_scale(const uint32_t* px, const RdoParams& p) {
  float m = rdo_channel_std_dev(px, 0);
  const float g = rdo_channel_std_dev(px, 1), b = rdo_channel_std_dev(px, 2), a = rdo_channel_std_dev(px, 3);
  m = m > g ? m : g; m = m > b ? m : b; m = m > a ? m : a;
  float yl = f_div(m, p.smooth_std_dev);
  yl = yl < 0.0f ? 0.0f : (yl > 1.0f ? 1.0f : yl);
  yl = f_mul(yl, yl);
  return f_add(p.smooth_scale, f_mul(f_sub(1.0f, p.smooth_scale), yl));  // lerp(a, b, t) = a + (b - a) * t
}

UB_FN float rdo_ms_err(uint64_t err) { return f_mul(u64_to_float(err), f_div(1.0f, 64.0f)); }

// ---- selector history: (first selector bit, first <= 64 selector bits) -> latest block index carrying them.
// Open addressing, never evicted inside a job range (the reference uses an unordered_map the same way).
struct RdoHistEntry { uint64_t sel; uint32_t ofs1; uint32_t index; };  // ofs1 = first_sel_bit + 1, 0 = empty slot

UB_FN uint32_t rdo_hash(uint64_t sel, uint32_t ofs) {
  uint64_t h = sel * 0x9E3779B97F4A7C15ull + (uint64_t)ofs * 0xC2B2AE3D27D4EB4Full;
  h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
  return (uint32_t)h;
}
// returns the stored index or -1
UB_FN int64_t rdo_hist_find(const RdoHistEntry* tab, uint32_t mask, uint32_t ofs, uint64_t sel) {
  uint32_t i = rdo_hash(sel, ofs) & mask;
  for (;;) {
    const RdoHistEntry e = tab[i];
    if (!e.ofs1) return -1;
    if (e.ofs1 == ofs + 1 && e.sel == sel) return (int64_t)e.index;
    i = (i + 1) & mask;
  }
}
UB_FN void rdo_hist_set(RdoHistEntry* tab, uint32_t mask, uint32_t ofs, uint64_t sel, uint32_t index) {
  uint32_t i = rdo_hash(sel, ofs) & mask;
  for (;;) {
    if (!tab[i].ofs1) { tab[i].sel = sel; tab[i].index = index; tab[i].ofs1 = ofs + 1; return; }
    if (tab[i].ofs1 == ofs + 1 && tab[i].sel == sel) { tab[i].index = index; return; }
    i = (i + 1) & mask;
  }
}

// ---- fast trial scoring.  A trial differs from the current block only in its weights, and both decodes (UASTC, and
// BC7 after transcoding) are per-texel functions of that texel's own weight(s): endpoints, partition, p-bits and the
// BC7 anchor normalisation do not depend on the other texels.  So the error of ANY trial is a sum of per-texel table
// entries  E[texel][plane][weight] = uastc_err + bc7_err of that texel (plane 1 = the dual-plane channel), built once
// per block by decoding the block with all weights set to each level in turn -- through the very same
// decode / transcode / BC7-decode code, so the tables are exact by construction.
struct RdoBlockInfo {
  uint8_t mode, wb, planes, npos;  // npos = anchor MSBs to re-insert into the selector field
  uint8_t pos[3];                  // their bit positions, ascending
  uint8_t valid;                   // 0 = the block does not unpack; solid blocks have mode == 8
};
const uint32_t RDO_TABLE_WORDS = 512;  // 16 texels x planes x levels <= 512 (mode 18: 16 x 1 x 32; dual plane: 16 x 2 x 4)

UB_FN uint32_t rdo_table_index(const RdoBlockInfo& f, uint32_t texel, uint32_t plane, uint32_t w) {
  return ((texel * f.planes + plane) << f.wb) + w;
}

UB_FN bool rdo_block_info(const Bits128& blk, RdoBlockInfo& f, UnpackedUastc& u) {
  memset(&f, 0, sizeof(f));
  if (!unpack_uastc(blk, u, false)) return false;
  f.valid = 1; f.mode = (uint8_t)u.mode;
  if (u.mode == UASTC_MODE_SOLID) return true;
  const uint32_t wb = md_weight_bits(u.desc), subsets = md_subsets(u.desc);
  f.wb = (uint8_t)wb; f.planes = (uint8_t)(1 + md_dual_plane(u.desc));
  if (f.planes == 2) { f.npos = 2; f.pos[0] = (uint8_t)(wb - 1); f.pos[1] = (uint8_t)(2 * wb - 1); }
  else {
    uint32_t a0 = u.anchors & 15, a1 = (u.anchors >> 4) & 15, a2 = (u.anchors >> 8) & 15, t;
    if (subsets == 1) { f.npos = 1; f.pos[0] = (uint8_t)(wb - 1); }
    else if (subsets == 2) {
      if (a0 > a1) { t = a0; a0 = a1; a1 = t; }
      f.npos = 2; f.pos[0] = (uint8_t)(a0 * wb + wb - 1); f.pos[1] = (uint8_t)(a1 * wb + wb - 1);
    } else {
      if (a0 > a1) { t = a0; a0 = a1; a1 = t; }
      if (a1 > a2) { t = a1; a1 = a2; a2 = t; }
      if (a0 > a1) { t = a0; a0 = a1; a1 = t; }
      f.npos = 3; f.pos[0] = (uint8_t)(a0 * wb + wb - 1); f.pos[1] = (uint8_t)(a1 * wb + wb - 1); f.pos[2] = (uint8_t)(a2 * wb + wb - 1);
    }
  }
  return true;
}

// Table column for one weight level: decode the block with every weight (both planes) set to `level`.
UB_BIG void rdo_fill_level(const uint32_t* px, const UnpackedUastc& u, const RdoBlockInfo& f, uint32_t level, uint32_t* table) {
  Candidate c;
  c.mode = u.mode; c.pattern = u.pattern; c.ccs = u.ccs;
  for (uint32_t i = 0; i < 18; i++) c.ep[i] = u.ep[i];
  for (uint32_t i = 0; i < 32; i++) c.w[i] = (uint8_t)level;
  uint32_t du[16], db[16];
  decode_uastc_pixels(c, du);
  UnpackedUastc t;
  candidate_to_unpacked(c, t);
  Bc7Build bb;
  uastc_to_bc7_build(t, bb);
  decode_bc7_build(bb, db);
  const uint32_t m1 = (f.planes == 2) ? (255u << (8 * u.ccs)) : 0u;  // the channel plane 1 drives
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t p = px[i];
    table[rdo_table_index(f, i, 0, level)] = sqdist4(p & ~m1, du[i] & ~m1) + sqdist4(p & ~m1, db[i] & ~m1);
    if (f.planes == 2) table[rdo_table_index(f, i, 1, level)] = sqdist4(p & m1, du[i] & m1) + sqdist4(p & m1, db[i] & m1);
  }
}

// (uastc_err + bc7_err) / 2 of the block whose selector field holds (lo, hi), from the block's table
UB_FN uint64_t rdo_fast_error(const uint32_t* table, const RdoBlockInfo& f, uint64_t lo, uint64_t hi) {
  Bits128 w;
  w.lo = lo; w.hi = hi;
  for (uint32_t k = 0; k < 3; k++) if (k < f.npos) w = bits_insert0(w, f.pos[k]);
  // entry e = texel * planes + plane sits at bit e * wb: peel the entries off a running 128-bit shift (branch-free; the walk's
  // trial loop spends most of its instructions here)
  uint32_t total = 0;
  const uint32_t wb = f.wb, mask = (1u << wb) - 1u;
  uint64_t a = w.lo, b = w.hi;
#pragma unroll
  for (uint32_t e = 0; e < 16; e++) {
    total += table[(e << wb) + ((uint32_t)a & mask)];
    a = (a >> wb) | (b << (64u - wb));
    b >>= wb;
  }
  if (f.planes == 2) {
#pragma unroll
    for (uint32_t e = 16; e < 32; e++) {
      total += table[(e << wb) + ((uint32_t)a & mask)];
      a = (a >> wb) | (b << (64u - wb));
      b >>= wb;
    }
  }
  return (uint64_t)(total >> 1);
}

// ---- the pieces of one step of the walk

// What the walk needs to know about the current block before the trials are judged (:3836-3928).
struct RdoCur {
  uint32_t mode, first, total;  // selector field of the block's mode
  uint64_t sel_lo;              // its own first <= 64 selector bits (the history key)
  float ms_err, rms_err, smooth;
  bool solid, skip;             // solid colour blocks are left alone; `skip` = error already above the threshold
};

UB_FN bool rdo_prepare(const uint32_t* px, const Bits128& blk, const RdoParams& p, RdoCur& c) {
  const uint32_t mode = UBT(uastc_huff_modes)[(uint32_t)blk.lo & 127u];
  if (mode >= UASTC_TOTAL_MODES) return false;
  c.mode = mode; c.solid = (mode == UASTC_MODE_SOLID); c.skip = false;
  c.first = UBT(rdo_sel_bits)[mode * 2]; c.total = UBT(rdo_sel_bits)[mode * 2 + 1];
  c.sel_lo = 0; c.ms_err = 0; c.rms_err = 0; c.smooth = 1.0f;
  if (c.solid) return true;
  c.smooth = rdo_smooth_scale(px, p);
  uint64_t err;
  if (!rdo_block_error(px, blk, err, nullptr)) return false;
  c.ms_err = rdo_ms_err(err);
  c.rms_err = f_sqrt(c.ms_err);
  uint64_t hi;
  rdo_get_sel(blk, c.first, c.total, c.sel_lo, hi);
  c.skip = (c.rms_err >= p.skip_rms);
  return true;
}

// Same, from the block's table (the drivers use this one; rdo_prepare is kept as the definition it must agree with).
UB_FN void rdo_prepare_fast(const uint32_t* px, const Bits128& blk, const RdoBlockInfo& f, const uint32_t* table, const RdoParams& p, RdoCur& c) {
  c.mode = f.mode; c.solid = (f.mode == UASTC_MODE_SOLID); c.skip = false;
  c.first = UBT(rdo_sel_bits)[c.mode * 2]; c.total = UBT(rdo_sel_bits)[c.mode * 2 + 1];
  c.sel_lo = 0; c.ms_err = 0; c.rms_err = 0; c.smooth = 1.0f;
  if (c.solid) return;
  c.smooth = rdo_smooth_scale(px, p);
  uint64_t hi;
  rdo_get_sel(blk, c.first, c.total, c.sel_lo, hi);
  c.ms_err = rdo_ms_err(rdo_fast_error(table, f, c.sel_lo, hi));
  c.rms_err = f_sqrt(c.ms_err);
  c.skip = (c.rms_err >= p.skip_rms);
}

// One trial from the table: prev's selector bits (slo, shi) spliced into the current block.
UB_FN bool rdo_trial_fast(const uint32_t* table, const RdoBlockInfo& f, uint64_t slo, uint64_t shi, const RdoCur& c,
                          uint32_t block_index, uint32_t match_index, const RdoParams& p, float& t) {
  const float ms = rdo_ms_err(rdo_fast_error(table, f, slo, shi));
  const float rms = f_sqrt(ms);
  if (rms > f_mul(c.rms_err, p.max_rms_ratio)) return false;
  const uint32_t bits = rdo_match_cost((block_index - match_index) * 16u);
  t = f_add(f_mul(ms, c.smooth), f_mul((float)(int)bits, p.lambda));
  return true;
}

// The chain-independent part of RdoCur, computed ahead of the walk (a block is only rewritten at its own step, so its
// own error, smoothness and selector bits are known before the walk reaches it).
struct RdoPre {
  float ms_err, rms_err, smooth;
  uint32_t skip;
  uint64_t sel_lo;
};
UB_FN RdoPre rdo_pre_of(const RdoCur& c) {
  RdoPre r;
  r.ms_err = c.ms_err; r.rms_err = c.rms_err; r.smooth = c.smooth; r.skip = c.skip ? 1u : 0u; r.sel_lo = c.sel_lo;
  return r;
}

// Modes whose selector field starts below bit 65 read bits that are ENDPOINT bits in a mode-0 block: a mode-0 block's
// endpoint re-fit may only be postponed past the walk if no such block follows it inside the window.
UB_FN bool rdo_mode_reads_mode0_endpoints(uint32_t mode) { return mode < UASTC_TOTAL_MODES && mode != UASTC_MODE_SOLID && UBT(rdo_sel_bits)[mode * 2] < 65; }

// cost of keeping the block as it is (:3917-3941): literal cost if its selector bits were never seen, else a match
UB_FN float rdo_keep_cost(const RdoCur& c, int64_t hist_index, uint32_t block_index, const RdoParams& p) {
  int bits;
  if (hist_index < 0) bits = (int)((c.total * p.literal_cost) / 100u);
  else bits = (int)rdo_match_cost((block_index - (uint32_t)hist_index) * 16u);
  return f_add(f_mul(c.ms_err, c.smooth), f_mul((float)bits, p.lambda));
}

// One trial (:3946-4012): splice prev's selector bits in, score, apply the RMS-increase gate, price the match.
// match_index = the history's block for prev's bits, or prev itself.  Returns false when the trial is rejected.
UB_FN bool rdo_trial(const uint32_t* px, const Bits128& cur, const Bits128& prev, const RdoCur& c, uint32_t block_index,
                     uint32_t match_index, const RdoParams& p, float& t, Bits128& trial) {
  trial = rdo_splice(cur, prev, c.first, c.total);
  uint64_t err;
  if (!rdo_block_error(px, trial, err, nullptr)) return false;
  const float ms = rdo_ms_err(err);
  const float rms = f_sqrt(ms);
  if (rms > f_mul(c.rms_err, p.max_rms_ratio)) return false;
  const uint32_t bits = rdo_match_cost((block_index - match_index) * 16u);
  t = f_add(f_mul(ms, c.smooth), f_mul((float)(int)bits, p.lambda));
  return true;
}

UB_FN uint64_t rgba_error16(const uint32_t* a, const uint32_t* b) {
  uint64_t e = 0;
  for (uint32_t i = 0; i < 16; i++) e += sqdist4(a[i], b[i]);
  return e;
}

// Mode-0 endpoint re-fit with the spliced selectors (:4026-4086).  Returns the block with the new endpoints when the
// UASTC error went down, else the block unchanged.  Hint bits are carried over untouched (they are recomputed by
// rdo_rehint afterwards; no later step of the walk reads them).
UB_BIG Bits128 rdo_refine_mode0(const uint32_t* px, const Bits128& blk) {
  UnpackedUastc u;
  if (!unpack_uastc(blk, u, false) || u.mode != 0) return blk;
  Candidate c0;
  unpacked_to_candidate(u, c0);
  uint32_t dec[16];
  decode_uastc_pixels(c0, dec);
  const uint64_t best_err = rgba_error16(px, dec);
  EncCfg cfg = enc_cfg_from_flags(0);
  cfg.uber_level = 0; cfg.ls_passes = 1;
  Candidate c1;
  mode_single_plane(0, 0, 0, px, cfg, c1, c0.w);
  decode_uastc_pixels(c1, dec);
  if (!(rgba_error16(px, dec) < best_err)) return blk;
  for (uint32_t i = 0; i < 32; i++) c1.w[i] = c0.w[i];  // selectors are the forced ones by construction
  BlockHints h;
  memset(&h, 0, sizeof(h));
  Bits128 nb = pack_uastc(c1, 0, h);
  const uint32_t d = UBT(uastc_mode_desc)[0];
  const uint64_t keep = lowmask64(md_huff_len(d) + md_hint_bits(d));  // mode code + hint field, < 64 bits
  nb.lo = (nb.lo & ~keep) | (blk.lo & keep);
  return nb;
}

// uastc_recompute_hints (:3648-3728) on a block whose selectors (and maybe endpoints) changed.
UB_BIG Bits128 rdo_rehint(const uint32_t* px, const Bits128& blk, uint32_t flags) {
  UnpackedUastc u;
  if (!unpack_uastc(blk, u, false) || u.mode == UASTC_MODE_SOLID) return blk;
  Candidate c;
  unpacked_to_candidate(u, c);
  uint32_t dec[16];
  decode_uastc_pixels(c, dec);
  // the reference takes the level from flags & 0xF here (not & 7 as encode_uastc does, :3671 vs :3180): with
  // cPackUASTCFavorUASTCError (8) set no level case matches, so the hint search runs with its default effort
  const uint32_t lvl = flags & 15u;
  EncCfg cfg = enc_cfg_from_flags(lvl > 4 ? 4u : lvl);
  cfg.level = lvl; cfg.flags = flags;
  BlockHints h;
  memset(&h, 0, sizeof(h));
  compute_hints(c, px, dec, cfg, h);
  return pack_uastc(c, 0, h);
}

// Job ranges of one image (:4117-4135): total_jobs ranges of n/total_jobs blocks plus a remainder range; one range
// when there is no pool, one job, or ranges would hold <= 8 blocks.
inline uint32_t rdo_job_ranges(uint32_t n, uint32_t total_jobs, uint32_t* blocks_per_job) {  // host-side planning
  const uint32_t per = total_jobs ? n / total_jobs : 0;
  if (total_jobs <= 1 || per <= 8) { *blocks_per_job = n; return n ? 1u : 0u; }
  *blocks_per_job = per;
  return (n + per - 1) / per;
}

}  // namespace ub
