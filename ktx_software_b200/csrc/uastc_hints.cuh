// uastc_hints.cuh -- transcoding hints stored in a UASTC block: BC1 hints, ETC1 flip/diff/intensity/bias, ETC2 EAC
// table+multiplier, and the solid-colour ETC1 encoding.  Restates, from the reference
// external/basis_universal/encoder/basisu_uastc_enc.cpp: compute_etc1_hints :2715-3011 (+ pack_etc1_estimate_flipped :2669,
// rgb_to_y_cb_cr / color_diff / gray_distance2 :2638-2667), uastc_pack_eac_a8 :3020-3105, compute_bc1_hints :2536-2629;
// from encoder/basisu_etc.cpp: pack_etc1_block_solid_color :181-262; from transcoder/basisu_transcoder.cpp:
// apply_etc1_bias :16182, encode_bc1 :17682-17860 (+ bc1_find_sels :17492, LS :17557, solid :17634),
// transcode_uastc_to_bc1_hint0/1 :18237/:18335; and unpack_bc1 (encoder/basisu_gpu_texture.cpp:94).
#pragma once
#include "uastc_modes.cuh"

namespace ub {

// ---------------------------------------------------------------------------------------------------------
// solid colour -> ETC1 (diff, inten, selector, r, g, b)
UB_BIG void hints_solid_etc1(uint32_t rgba, uint32_t& diff, uint32_t& inten, uint32_t& sel, uint32_t& r, uint32_t& g, uint32_t& b) {
  const uint32_t next_comp[4] = {1, 2, 0, 1};
  uint32_t col[3] = {ch(rgba, 0), ch(rgba, 1), ch(rgba, 2)};
  uint32_t best_error = 0xFFFFFFFFu, best_i = 0, best_x = 0, best_c1 = 0, best_c2 = 0;
  bool done = false;
  for (uint32_t i = 0; i < 3 && !done; i++) {
    const uint32_t c1 = col[next_comp[i]], c2 = col[next_comp[i + 1]];
    for (int delta = -1; delta <= 1 && !done; delta++) {
      const int cpd = clampi((int)col[i] + delta, 0, 255);
      const uint16_t* p = UBT(etc1_solid_cfg) + UBT(etc1_solid_cfg_ofs)[cpd];
      do {
        const uint32_t x = *p++;
        const uint16_t* inv = UBT(etc1_inverse) + (x & 0xFFu) * 256;
        const uint32_t p1 = inv[c1], p2 = inv[c2];
        const int d0 = cpd - (int)col[i];
        const uint32_t trial = (uint32_t)(d0 * d0) + (p1 >> 8) * (p1 >> 8) + (p2 >> 8) * (p2 >> 8);
        if (trial < best_error) {
          best_error = trial; best_x = x; best_c1 = p1 & 0xFFu; best_c2 = p2 & 0xFFu; best_i = i;
          if (!best_error) { done = true; break; }
        }
      } while (*p != 0xFFFF);
    }
  }
  diff = best_x & 1u; inten = (best_x >> 1) & 7u; sel = (best_x >> 4) & 3u;
  uint32_t out[3];
  out[best_i] = (best_x >> 8) & 255u;
  out[next_comp[best_i]] = best_c1;
  out[next_comp[best_i + 1]] = best_c2;
  r = out[0]; g = out[1]; b = out[2];
}

// ---------------------------------------------------------------------------------------------------------
// ETC1 hints
struct YCbCr { int y, cb, cr; };
UB_FN YCbCr to_ycbcr(uint32_t c) {
  YCbCr o;
  o.y = (int)ch(c, 0) * 54 + (int)ch(c, 1) * 183 + (int)ch(c, 2) * 19;
  o.cb = ((int)ch(c, 2) << 8) - o.y;
  o.cr = ((int)ch(c, 0) << 8) - o.y;
  return o;
}
UB_FN uint64_t ycc_diff(const YCbCr& a, const YCbCr& b) {
  const int64_t dy = a.y - b.y, dcb = a.cb - b.cb, dcr = a.cr - b.cr;
  return (uint64_t)(dy * dy * 4 + dcr * dcr + dcb * dcb);
}
UB_FN float u64_to_float_h(uint64_t v) {
#ifdef __CUDA_ARCH__
  return __ull2float_rn(v);
#else
  return (float)v;
#endif
}
UB_FN int clamp255(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }

UB_FN int gray_distance2(uint32_t c, int r, int g, int b) {
  const int gd = (((int)ch(c, 0) - r) + ((int)ch(c, 1) - g) + ((int)ch(c, 2) - b) + 1) / 3;
  const int dr = (int)ch(c, 0) - clamp255(r + gd), dg = (int)ch(c, 1) - clamp255(g + gd), db = (int)ch(c, 2) - clamp255(b + gd);
  return dr * dr + dg * dg + db * db;
}

UB_FN bool etc1_estimate_flipped(const uint32_t* px) {
  int upper[3], lower[3], left[3], right[3];
#pragma unroll
  for (uint32_t c = 0; c < 3; c++) {
    int s00 = 0, s10 = 0, s01 = 0, s11 = 0;
    for (uint32_t y = 0; y < 2; y++)
      for (uint32_t x = 0; x < 2; x++) {
        s00 += (int)ch(px[x + y * 4], c); s10 += (int)ch(px[(x + 2) + y * 4], c);
        s01 += (int)ch(px[x + (y + 2) * 4], c); s11 += (int)ch(px[(x + 2) + (y + 2) * 4], c);
      }
    upper[c] = (s00 + s10 + 4) / 8; lower[c] = (s01 + s11 + 4) / 8; left[c] = (s00 + s01 + 4) / 8; right[c] = (s10 + s11 + 4) / 8;
  }
  int ug = 0, lg = 0, leg = 0, rg = 0;
#pragma unroll 1
  for (uint32_t i = 0; i < 4; i++)
#pragma unroll 1
    for (uint32_t j = 0; j < 2; j++) {
      ug += gray_distance2(px[i + j * 4], upper[0], upper[1], upper[2]);
      lg += gray_distance2(px[i + (2 + j) * 4], lower[0], lower[1], lower[2]);
      leg += gray_distance2(px[j + i * 4], left[0], left[1], left[2]);
      rg += gray_distance2(px[(2 + j) + i * 4], right[0], right[1], right[2]);
    }
  return (ug + lg) < (leg + rg);
}

// apply_etc1_bias (XC:16182) as a table: the per-channel delta of every bias mode, [bias][sub-block][channel].
UB_TABLE(int8_t, etc1_bias_delta, 192) = {
  -1,-1,-1,-1,-1,-1,0,-1,-1,0,-1,-1,-1,0,0,0,0,0,-1,0,-1,-1,0,-1,0,0,-1,0,0,-1,0,-1,0,0,0,0,0,0,-1,0,0,0,1,0,0,0,0,0,
  0,0,0,0,0,1,-1,-1,0,-1,-1,0,-2,-2,-2,-2,-2,-2,0,1,0,0,0,0,-1,0,0,-1,0,0,0,0,0,0,0,0,1,0,0,1,0,0,0,0,1,0,0,0,
  0,1,0,0,1,0,1,1,0,1,1,0,0,0,0,-1,0,0,0,0,0,0,-1,0,0,0,0,0,0,-1,0,0,0,1,0,0,0,0,1,0,0,1,1,0,1,1,0,1,
  0,0,0,0,1,0,0,1,1,0,1,1,1,1,1,1,1,1,-1,-1,-1,0,0,0,1,1,1,-1,-1,-1,0,0,0,1,1,1,0,0,0,-1,-1,-1,1,1,1,0,0,0};
UB_TABLE(uint8_t, etc1_sorted_bias, 32) = {13, 0, 22, 29, 27, 12, 26, 9, 30, 31, 8, 10, 25, 2, 23, 5, 15, 7, 3, 11, 6, 17, 28, 18, 1, 19, 20, 21, 24, 4, 14, 16};

// Memo slot of bias trial `bi` for sub-block s: trials whose bias moves the sub-block's base colour by the same
// per-channel steps share a slot.  [sorted order (levels <= 3) | plain order (level 4)][sub-block][bi]
UB_TABLE(uint8_t, etc1_bias_slot, 128) = {
  0,1,2,0,1,3,4,5,0,4,0,6,7,3,8,9,2,10,11,12,13,14,4,0,15,0,0,0,0,13,10,12,0,1,2,3,0,4,3,5,1,0,2,6,7,0,8,0,0,0,9,0,0,10,1,4,11,12,13,14,15,13,14,15,
  0,1,2,3,4,5,4,6,7,8,9,10,2,7,6,11,10,12,7,7,7,7,11,13,7,14,15,0,15,7,7,15,0,1,2,3,4,2,2,2,5,6,7,2,8,2,9,2,10,11,8,12,4,9,5,13,10,14,15,2,0,15,0,2};

UB_FN int etc1_bias_comp(int v, uint32_t bias, int limit, uint32_t subblock, uint32_t c) {  // apply_etc1_bias, one channel
  const int delta = UBT(etc1_bias_delta)[bias * 6 + subblock * 3 + c];
  if (v == 0) { if (delta == -2) v += 3; else v += delta + 1; }
  else if (v == limit) v += (delta - 1);
  else { v += delta; if (v < 0 || v > limit) v = (v - delta) - delta; }
  return v;
}

// texel index j (0..7) of sub-block `subset` under `flip`
UB_FN uint32_t etc1_texel(uint32_t flip, uint32_t subset, uint32_t j) {
  return flip ? ((subset * 2 + (j >> 2)) * 4 + (j & 3u)) : ((j >> 1) * 4 + subset * 2 + (j & 1u));
}

// YCbCr of a texel, luma pre-doubled: the metric is 4*dy^2 + dcr^2 + dcb^2 == (2dy)^2 + dcr^2 + dcb^2 (color_diff :2649)
struct Ycc2 { int y2, cb, cr; };
UB_FN Ycc2 to_ycc2(uint32_t c) {
  const YCbCr t = to_ycbcr(c);
  Ycc2 o; o.y2 = t.y * 2; o.cb = t.cb; o.cr = t.cr;
  return o;
}
UB_FN uint64_t sq64(int d) { return (uint64_t)((int64_t)d * (int64_t)d); }
UB_FN uint64_t chroma_d(const Ycc2& a, const Ycc2& b) { return sq64(a.cr - b.cr) + sq64(a.cb - b.cb); }

// Moments of the chroma of 8 texels: sum_j chroma_d(b, texel_j) == 8 (b.cr^2 + b.cb^2) - 2 (b.cr S1cr + b.cb S1cb) + S2
struct EtcMoments {
  int64_t s1cr, s1cb, s2;
  UB_MFN void clear() { s1cr = 0; s1cb = 0; s2 = 0; }
  UB_MFN void add(const Ycc2& t) { s1cr += t.cr; s1cb += t.cb; s2 += (int64_t)t.cr * t.cr + (int64_t)t.cb * t.cb; }
  UB_MFN uint64_t chroma_sum(const Ycc2& b) const {
    return (uint64_t)(8 * ((int64_t)b.cr * b.cr + (int64_t)b.cb * b.cb) - 2 * ((int64_t)b.cr * s1cr + (int64_t)b.cb * s1cb) + s2);
  }
};

// One ETC1 sub-block's view of a block under a given flip: the 8 decoded texels (selectors are picked on these) and the
// 8 source texels (the reported error is against these), plus what the search needs of them in aggregate.
struct Etc1Sub {
  int dy[8], ndcr[8], ndcb[8];  // decoded texels: luma (not doubled), chroma NEGATED (the form the clamping-table scan multiplies)
  int sy2[8], scr[8], scb[8];   // source texels: doubled luma, chroma
  uint32_t drgb[8];             // decoded texels, packed
  EtcMoments dmom, smom;
  int64_t dnorm;                // sum over the decoded texels of 4 y^2 + cr^2 + cb^2
  int mn[3], mx[3];             // per-channel range of the decoded texels
  uint32_t sum[3];              // per-channel sum of the decoded texels
  uint32_t pad_;                // keeps the two sub-blocks of a pair in different shared-memory banks
};

// Search limits that depend only on mode / level / flags.
struct Etc1Plan { uint32_t last_bias, last_individ; bool has_bias, use_table, flip_estimate; };
UB_FN Etc1Plan etc1_plan(uint32_t mode, uint32_t level, uint32_t flags) {
  const bool faster = (flags & 64u) != 0, fastest = (flags & 128u) != 0;
  Etc1Plan p;
  p.has_bias = UBT(uastc_mode_has_etc1_bias)[mode] != 0;
  p.last_bias = 1; p.use_table = false;
  p.flip_estimate = (level <= 1) || faster || fastest;
  if (p.has_bias) {
    p.use_table = (level <= 3);
    switch (level) {
      case 0: p.last_bias = fastest ? 1 : (faster ? 1 : 2); break;
      case 1: p.last_bias = fastest ? 1 : (faster ? 3 : 5); break;
      case 2: p.last_bias = fastest ? 1 : (faster ? 10 : 20); break;
      case 3: p.last_bias = fastest ? 1 : (faster ? 16 : 32); break;
      default: p.last_bias = 32; break;
    }
  }
  p.last_individ = (flags & 256u) ? 1u : 2u;
  return p;
}

// Base colours (8-bit, per sub-block) of bias trial `bi` in individual (4:4:4 x2) or differential (5:5:5 + 3:3:3) mode.
UB_FN void etc1_base_colors(const uint32_t* sum0, const uint32_t* sum1, const Etc1Plan& pl, uint32_t individ, uint32_t bi, int (*base8)[3],
                            uint32_t& bias) {
  const uint32_t mul = individ ? 15u : 31u;
  bias = pl.use_table ? UBT(etc1_sorted_bias)[bi] : bi;
#pragma unroll
  for (uint32_t c = 0; c < 3; c++) {
    const int u0 = (int)((sum0[c] * mul + 1020u) / (8u * 255u)), u1 = (int)((sum1[c] * mul + 1020u) / (8u * 255u));
    const int b0 = pl.has_bias ? etc1_bias_comp(u0, bias, (int)mul, 0, c) : u0;
    const int b1 = pl.has_bias ? etc1_bias_comp(u1, bias, (int)mul, 1, c) : u1;
    if (individ) {
      base8[0][c] = (b0 << 4) | b0;
      base8[1][c] = (b1 << 4) | b1;
    } else {
      const int d = clampi(b1 - b0, -4, 3);
      const int c1 = clampi(b0 + d, 0, 31);
      base8[0][c] = (b0 << 3) | (b0 >> 2);
      base8[1][c] = (c1 << 3) | (c1 >> 2);
    }
  }
}

// One colour P of an ETC1 sub-block (base + modifier, clamped) prepared for distance tests against many texels T, in the
// weighted (2y, cr, cb) space: |P - T|^2 - |T|^2 = |P|^2 - 2 P.T.  The scan wants, per texel, the entry of least distance, the
// lowest entry index among equals: all of it times four with the entry index in the two low bits makes that ONE 64-bit minimum
// (4 g + k orders like (g, k)), and the index comes back out of the winner's low bits.
struct EtcEntry {
  int64_t n4k;      // 4 (4 y^2 + cr^2 + cb^2) + k
  int a4, b4, c4;   // 32 y, 8 cr, 8 cb: multiplied by the NEGATED texel (-ty, -tcr, -tcb)
  // c + a * b, 32 x 32 -> 64 signed: one IMAD.WIDE.  (Left to the compiler, the product with a4 -- known to be >= 0 -- became an
  // unsigned wide multiply plus a three-instruction sign fix-up.)
  static UB_MFN int64_t mad_wide(int a, int b, int64_t c) {
#ifdef __CUDA_ARCH__
    int64_t d;
    asm("mad.wide.s32 %0, %1, %2, %3;" : "=l"(d) : "r"(a), "r"(b), "l"(c));
    return d;
#else
    return c + (int64_t)a * b;
#endif
  }
  UB_MFN int64_t d4(int nty, int ntcr, int ntcb) const { return mad_wide(c4, ntcb, mad_wide(b4, ntcr, mad_wide(a4, nty, n4k))); }
};
UB_FN void etc1_table_entries(const int* base8, uint32_t inten, EtcEntry* e) {
#pragma unroll
  for (uint32_t k = 0; k < 4; k++) {
    const int m = UBT(etc1_inten)[inten * 4 + k];
    const int r = clamp255(base8[0] + m), g = clamp255(base8[1] + m), b = clamp255(base8[2] + m);
    const int y = r * 54 + g * 183 + b * 19, cr = (r << 8) - y, cb = (b << 8) - y;
    e[k].n4k = 4 * (4 * (int64_t)y * y + (int64_t)cr * cr + (int64_t)cb * cb) + (int64_t)k;
    e[k].a4 = 32 * y; e[k].b4 = 8 * cr; e[k].c4 = 8 * cb;
  }
}

// Texel j of sub-block s under `flip` into the sub-block view; etc1_sub_finish then derives the aggregates.
UB_FN void etc1_sub_set_texel(Etc1Sub& sb, uint32_t j, uint32_t dec_px, uint32_t src_px) {
  const Ycc2 d = to_ycc2(dec_px), q = to_ycc2(src_px);
  sb.dy[j] = d.y2 >> 1; sb.ndcr[j] = -d.cr; sb.ndcb[j] = -d.cb;
  sb.sy2[j] = q.y2; sb.scr[j] = q.cr; sb.scb[j] = q.cb;
  sb.drgb[j] = dec_px;
}
UB_FN void etc1_sub_finish(Etc1Sub& sb) {
#pragma unroll
  for (uint32_t c = 0; c < 3; c++) { sb.sum[c] = 0; sb.mn[c] = 255; sb.mx[c] = 0; }
  sb.dmom.clear(); sb.smom.clear(); sb.dnorm = 0;
#pragma unroll 1
  for (uint32_t j = 0; j < 8; j++) {
    const uint32_t p = sb.drgb[j];
#pragma unroll
    for (uint32_t c = 0; c < 3; c++) {
      const int v = (int)ch(p, c);
      sb.sum[c] += (uint32_t)v;
      sb.mn[c] = v < sb.mn[c] ? v : sb.mn[c];
      sb.mx[c] = v > sb.mx[c] ? v : sb.mx[c];
    }
    Ycc2 d; d.y2 = 0; d.cr = -sb.ndcr[j]; d.cb = -sb.ndcb[j];
    Ycc2 q; q.y2 = 0; q.cr = sb.scr[j]; q.cb = sb.scb[j];
    sb.dmom.add(d); sb.smom.add(q);
    sb.dnorm += 4 * (int64_t)sb.dy[j] * sb.dy[j] + (int64_t)d.cr * d.cr + (int64_t)d.cb * d.cb;
  }
}

// The two sub-block views of one flip.
UB_FN void etc1_hints_setup(Etc1Sub* sub, uint32_t flip, const uint32_t* px, const uint32_t* dec) {
#pragma unroll 1
  for (uint32_t s = 0; s < 2; s++) {
#pragma unroll 1
    for (uint32_t j = 0; j < 8; j++) {
      const uint32_t t = etc1_texel(flip, s, j);
      etc1_sub_set_texel(sub[s], j, dec[t], px[t]);
    }
    etc1_sub_finish(sub[s]);
  }
}

// All bias trials of one (flip, colour mode): improves (best_err, o_*) where a trial is strictly better, in trial order.
//
// An evaluation = one (sub-block, base colour): scan 2 / 4 / 8 intensity tables (by the range of the texels around the base
// colour) for the lowest error against the DECODED texels -- in table order, first of equals, and in the non-flipped layout the
// scan stops at the first table that does not improve -- then report the chosen table's error against the SOURCE texels.  Its
// result depends only on (flip, sub-block, base colour): bias trials that step a sub-block's base colour alike share it.
//
// The scan of a table whose modifiers push the base colour past 0 or 255 (4 clamped entries x 8 texels of 64-bit distance
// arithmetic, ~8x a plain table) is most of the search.  Run trial by trial, a warp paid for it whenever ANY lane's current table
// clamped while the other lanes waited (13 of 32 lanes active on those lines).  So the search runs in four passes:
//   1  per trial: base colours, the distinct evaluations listed (memo by slot, as the reference's trial order finds them), their
//      plain tables scanned (the tables' larger modifiers grow with the table index, so an evaluation's tables are plain below
//      n_plain and clamping from there on);
//   2  the clamping tables: every lane walks ITS list of open evaluations, one clamping table per step, the lanes of the warp
//      stepping together -- no lane waits for another's table, only for the lane with the most of them at the very end;
//   3  per evaluation: the chosen table's error against the source texels (a clamping table reuses the entry picks recorded
//      when it was scored);
//   4  per trial, in trial order: the sum of its two evaluations against the best so far.
// `lanes` (device): the lanes of the warp that run this call together -- they step through pass 2 together; 0 = the caller's
// lanes are not known to be together (thread = block callers).
// index of the lowest set bit (x != 0)
UB_FN uint32_t ctz64(uint64_t x) {
#ifdef __CUDA_ARCH__
  return (uint32_t)__ffsll((long long)x) - 1u;
#else
  return (uint32_t)__builtin_ctzll(x);
#endif
}

#ifndef ETC1_CLAMP_UNROLL
#define ETC1_CLAMP_UNROLL 4   // texels per iteration of the clamping-table scan: 1 / 2 / 4 / 8 measured 24.3 / 23.6 / 23.35 / 23.6 ms (4096^2 img_mix level 2)
#endif
constexpr int ETC1_CLAMP_UNROLL_N = ETC1_CLAMP_UNROLL;
struct Etc1Eval {
  uint32_t key;     // base colour, r | g << 8 | b << 16
  uint32_t st;      // sub-block | limit << 4 | n_plain << 8 | next table << 12 (== limit: scan over) | best table << 16
  uint64_t best;    // lowest table error so far; after pass 3 the error against the source texels
  uint32_t picks;   // entry of every texel (2 bits each) under the best table, when that is a clamping one
  uint32_t pad_;
};
UB_FN void etc1_hints_search(const Etc1Plan& pl, const Etc1Sub* sub, uint32_t flip, uint32_t individ, uint32_t level, uint64_t& best_err,
                             uint32_t& o_flip, uint32_t& o_diff, uint32_t& o_i0, uint32_t& o_i1, uint32_t& o_bias, uint32_t lanes = 0) {
  Etc1Eval ev[64];          // at most two new evaluations per trial, at most 32 trials
  uint8_t slot_ev[2][16];   // the evaluation a memo slot (etc1_bias_slot) holds; differential mode couples the second sub-block's
                            // colour to the first's, so a slot can meet a second colour: that is a new evaluation
  uint8_t trial_ev[32][2];
  uint32_t n_ev = 0;
  uint64_t open = 0;        // evaluations with clamping tables still to scan (bit = index in ev): the cursor of pass 2 lives in registers
#pragma unroll 1
  for (uint32_t i = 0; i < 16; i++) { slot_ev[0][i] = 0xFF; slot_ev[1][i] = 0xFF; }
  // ---- pass 1
#pragma unroll 1
  for (uint32_t bi = 0; bi < pl.last_bias; bi++) {
    int base8[2][3];
    uint32_t bias;
    etc1_base_colors(sub[0].sum, sub[1].sum, pl, individ, bi, base8, bias);
#pragma unroll 1
    for (uint32_t s = 0; s < 2; s++) {
      const int* b8 = s ? base8[1] : base8[0];
      const uint32_t key = (uint32_t)b8[0] | ((uint32_t)b8[1] << 8) | ((uint32_t)b8[2] << 16);
      const uint32_t slot = UBT(etc1_bias_slot)[(pl.use_table ? 0u : 64u) + s * 32 + bi];
      const uint32_t held = slot_ev[s][slot];
      if (held != 0xFFu && ev[held].key == key) { trial_ev[bi][s] = (uint8_t)held; continue; }
      const Etc1Sub& sb = sub[s];
      int range = 0;
#pragma unroll
      for (uint32_t c = 0; c < 3; c++) {
        int a = sb.mx[c] - b8[c]; if (a < 0) a = -a;
        int b = b8[c] - sb.mn[c]; if (b < 0) b = -b;
        if (a > range) range = a;
        if (b > range) range = b;
      }
      const uint32_t limit = (level == 4) ? 8u : ((range > 51) ? 8u : (range >= 7 ? 4u : 2u));
      const Ycc2 by = to_ycc2(key);
      const int yb = by.y2 >> 1;
      const int bmin = b8[0] < b8[1] ? (b8[0] < b8[2] ? b8[0] : b8[2]) : (b8[1] < b8[2] ? b8[1] : b8[2]);
      const int bmax = b8[0] > b8[1] ? (b8[0] > b8[2] ? b8[0] : b8[2]) : (b8[1] > b8[2] ? b8[1] : b8[2]);
      const int room = bmin < 255 - bmax ? bmin : 255 - bmax;   // a table is plain while its larger modifier fits
      // larger modifiers of the tables: 8, 17, 29, 42, 60, 80, 106, 183
      uint32_t n_plain = (uint32_t)((room >= 8) + (room >= 17) + (room >= 29) + (room >= 42) + (room >= 60) + (room >= 80) + (room >= 106) + (room >= 183));
      n_plain = n_plain < limit ? n_plain : limit;
      const uint64_t ksum = sb.dmom.chroma_sum(by);
      uint64_t best_sub = ~0ull;
      uint32_t bi_s = 0, it = 0;
#pragma unroll 1
      while (it < n_plain) {
        // No entry clamps: all four share the base chroma and sit at luma yb + 256*m.  The nearest entry to a
        // texel follows from |dy| against the midpoint 128*(m0+m1); the squared distance fits 32 bits.
        const int m0 = UBT(etc1_inten)[it * 4 + 2], m1 = UBT(etc1_inten)[it * 4 + 3];  // table = {-m1, -m0, m0, m1}
        const int mid = 128 * (m0 + m1), o0 = 256 * m0, o1 = 256 * m1;
        uint64_t total = 0;
#pragma unroll
        for (uint32_t j = 0; j < 8; j++) {
          int d = sb.dy[j] - yb; d = d < 0 ? -d : d;
          const int x = d - (d < mid ? o0 : o1);
          total += (uint32_t)(x * x);
        }
        total = total * 4 + ksum;
        // the reference stops scanning tables at the first non-improving one only in the non-flipped layout
        if (!flip && total >= best_sub) { it = limit; break; }
        if (total < best_sub) { best_sub = total; bi_s = it; }
        it++;
      }
      Etc1Eval& e = ev[n_ev];
      e.key = key; e.st = s | (limit << 4) | (n_plain << 8) | (it << 12) | (bi_s << 16); e.best = best_sub; e.picks = 0; e.pad_ = 0;
      slot_ev[s][slot] = (uint8_t)n_ev; trial_ev[bi][s] = (uint8_t)n_ev;
      if (it < limit) open |= 1ull << n_ev;
      n_ev++;
    }
  }
  // ---- pass 2: one clamping table per step
  uint64_t by_picks = 0;    // evaluations whose best table is a clamping one
  {
    for (;;) {
      const bool done = open == 0;
#ifdef __CUDA_ARCH__
      if (lanes) { if (!__any_sync(lanes, !done)) break; }
      else
#endif
      if (done) break;
      if (!done) {
        // |P - T|^2 = |P|^2 - 2 P.T + |T|^2 in the weighted (2y, cr, cb) space: per entry the constant |P|^2 and three
        // coefficients, per texel three multiply-adds and a compare; sum |T|^2 is a sub-block constant (EtcEntry)
        const uint32_t cur = ctz64(open);
        Etc1Eval& e = ev[cur];
        const uint32_t st = e.st, it = (st >> 12) & 15u, limit = (st >> 4) & 15u;
        const Etc1Sub& sb = sub[st & 1u];
        const int b8[3] = {(int)(e.key & 255u), (int)((e.key >> 8) & 255u), (int)((e.key >> 16) & 255u)};
        EtcEntry en[4];
        etc1_table_entries(b8, it, en);
        int64_t acc4 = 0;
        uint32_t picks = 0;
#pragma unroll ETC1_CLAMP_UNROLL_N
        for (uint32_t j = 0; j < 8; j++) {
          const int nty = -sb.dy[j], ntcr = sb.ndcr[j], ntcb = sb.ndcb[j];
          int64_t m = en[0].d4(nty, ntcr, ntcb);
#pragma unroll
          for (uint32_t k = 1; k < 4; k++) {
            const int64_t g = en[k].d4(nty, ntcr, ntcb);
            m = g < m ? g : m;
          }
          acc4 += m;
          picks |= ((uint32_t)m & 3u) << (2 * j);
        }
        uint32_t ks = (picks & 0x3333u) + ((picks >> 2) & 0x3333u);   // sum of the eight entry indices
        ks = (ks & 0x0F0Fu) + ((ks >> 4) & 0x0F0Fu);
        ks = (ks & 0xFFu) + (ks >> 8);
        const uint64_t total = (uint64_t)(sb.dnorm + ((acc4 - (int64_t)ks) >> 2));   // acc4 - ks = 4 x the sum of the distances, exactly
        uint32_t next = it + 1, st2 = st;
        if (!flip && total >= e.best) next = limit;
        else if (total < e.best) { e.best = total; e.picks = picks; st2 = (st & 0xFFFFu) | (it << 16); by_picks |= 1ull << cur; }
        e.st = (st2 & ~0xF000u) | (next << 12);
        if (next >= limit) open &= ~(1ull << cur);
      }
    }
  }
  // ---- pass 3: error of the chosen table against the ORIGINAL texels (selectors picked on the decoded ones; equal
  // distances keep the lowest entry index).  Two loops, each over a lane's own evaluations of one kind, so the bodies run converged.
  {
    uint64_t todo = (n_ev >= 64u ? ~0ull : ((1ull << n_ev) - 1ull)) & ~by_picks;   // best table is a plain one
#pragma unroll 1
    while (todo) {
      const uint32_t i = ctz64(todo);
      todo &= todo - 1;
      Etc1Eval& e = ev[i];
      const uint32_t st = e.st, bi_s = st >> 16;
      const Etc1Sub& sb = sub[st & 1u];
      const Ycc2 by = to_ycc2(e.key);
      const int yb = by.y2 >> 1;
      const int m0 = UBT(etc1_inten)[bi_s * 4 + 2], m1 = UBT(etc1_inten)[bi_s * 4 + 3];
      const int mid = 128 * (m0 + m1);
      uint64_t err_s = 0;
#pragma unroll
      for (uint32_t j = 0; j < 8; j++) {
        const int sd = sb.dy[j] - yb, d = sd < 0 ? -sd : sd;
        // entries by luma: k0 = -m1, k1 = -m0, k2 = +m0, k3 = +m1
        const int m = (sd <= 0) ? ((d >= mid) ? -m1 : -m0) : ((d > mid) ? m1 : m0);
        err_s += sq64(sb.sy2[j] - (by.y2 + 512 * m));
      }
      e.best = err_s + sb.smom.chroma_sum(by);
    }
    todo = by_picks;   // best table is a clamping one: the entry of every texel was recorded when it was scored
#pragma unroll 1
    while (todo) {
      const uint32_t i = ctz64(todo);
      todo &= todo - 1;
      Etc1Eval& e = ev[i];
      const uint32_t st = e.st, bi_s = st >> 16, picks = e.picks;
      const Etc1Sub& sb = sub[st & 1u];
      const int b8[3] = {(int)(e.key & 255u), (int)((e.key >> 8) & 255u), (int)((e.key >> 16) & 255u)};
      uint64_t err_s = 0;
#pragma unroll 2
      for (uint32_t j = 0; j < 8; j++) {
        const int m = UBT(etc1_inten)[bi_s * 4 + ((picks >> (2 * j)) & 3u)];
        const int r = clamp255(b8[0] + m), g = clamp255(b8[1] + m), b = clamp255(b8[2] + m);
        const int y = r * 54 + g * 183 + b * 19, cr = (r << 8) - y, cb = (b << 8) - y;
        const int ey = y - (sb.sy2[j] >> 1), ecr = cr - sb.scr[j], ecb = cb - sb.scb[j];
        err_s += (uint64_t)(4 * (int64_t)ey * ey + (int64_t)ecr * ecr + (int64_t)ecb * ecb);
      }
      e.best = err_s;
    }
  }
  // ---- pass 4
#pragma unroll 1
  for (uint32_t bi = 0; bi < pl.last_bias; bi++) {
    const Etc1Eval& e0 = ev[trial_ev[bi][0]];
    const Etc1Eval& e1 = ev[trial_ev[bi][1]];
    const uint64_t err = e0.best + e1.best;
    if (err < best_err) {
      best_err = err; o_flip = flip; o_diff = individ ? 0u : 1u; o_i0 = e0.st >> 16; o_i1 = e1.st >> 16;
      o_bias = pl.use_table ? UBT(etc1_sorted_bias)[bi] : bi;
    }
  }
}

// Which (flip, colour mode) pairs the search visits, in its order: flips [first, last), colour modes [0, last_individ).
UB_FN void etc1_hints_range(const Etc1Plan& pl, uint32_t flags, const uint32_t* dec, uint32_t& first_flip, uint32_t& last_flip) {
  first_flip = 0; last_flip = 2;
  if (flags & 256u) last_flip = 1;
  else if (pl.flip_estimate) { if (etc1_estimate_flipped(dec)) first_flip = 1; last_flip = first_flip + 1; }
}

// One thread, one block.
UB_BIG void compute_etc1_hints(uint32_t mode, const uint32_t* px, const uint32_t* dec, uint32_t level, uint32_t flags,
                              uint32_t& o_flip, uint32_t& o_diff, uint32_t& o_i0, uint32_t& o_i1, uint32_t& o_bias) {
  const Etc1Plan pl = etc1_plan(mode, level, flags);
  uint32_t first_flip, last_flip;
  etc1_hints_range(pl, flags, dec, first_flip, last_flip);
  uint64_t best_err = ~0ull;
  o_flip = 0; o_diff = 0; o_i0 = 0; o_i1 = 0; o_bias = 0;
  Etc1Sub sub[2];
#pragma unroll 1
  for (uint32_t flip = first_flip; flip < last_flip; flip++) {
    etc1_hints_setup(sub, flip, px, dec);
#pragma unroll 1
    for (uint32_t individ = 0; individ < pl.last_individ; individ++)
      etc1_hints_search(pl, sub, flip, individ, level, best_err, o_flip, o_diff, o_i0, o_i1, o_bias);
  }
}

// The same search as independent jobs (uastc_jobs.cuh): job (fslot, individ) of a block returns the best trial of ITS
// (flip, colour mode) pair; etc1_hints_combine over the jobs in (fslot, individ) order with a strict compare picks what the
// loop above picks.  fslot counts from the block's first flip.  A job outside the block's range returns err = ~0.
const uint32_t ETC1_JOB_BOTH_MODES = 0xFFu;   // `individ` of a job that runs the colour modes one after the other (shares the sub-block set-up)
struct Etc1JobOut { uint64_t err; uint32_t hint; uint32_t pad_; };   // hint = flip | diff << 1 | inten0 << 2 | inten1 << 5 | bias << 8
UB_FN uint32_t etc1_job_slots(uint32_t level, uint32_t flags) {   // bit 0..1: flip slots, bit 2..3: colour modes (uniform over a launch)
  const Etc1Plan pl = etc1_plan(0, level, flags);
  const uint32_t nf = ((flags & 256u) || pl.flip_estimate) ? 1u : 2u;
  return nf | (pl.last_individ << 2);
}
// Cost class of a block's ETC1 hint search, 0 (dearest) .. ETC1_COST_CLASSES - 1: a proxy, used only to put blocks of similar
// cost into one warp.  A sub-block evaluation scans 2 / 4 / 8 intensity tables (by the range of its texels around the base
// colour), and a table whose modifiers push the base colour past 0 or 255 takes the clamping path, ~8x the plain one; lanes on
// the plain path idle while one lane of the warp clamps.  The proxy counts, per sub-block view, tables and clamping tables for
// the view's average colour.
const uint32_t ETC1_COST_CLASSES = 12;
UB_FN uint32_t etc1_cost_class(uint32_t mode, const uint32_t* dec) {
  // per 2x2 quadrant and channel: sum, min, max; the four sub-block views (flip 0: left / right, flip 1: top / bottom) pair them up
  int qs[4][3], qmn[4][3], qmx[4][3];
#pragma unroll
  for (uint32_t q = 0; q < 4; q++) {
    const uint32_t t0 = (q >> 1) * 8 + (q & 1u) * 2;
#pragma unroll
    for (uint32_t c = 0; c < 3; c++) {
      const int a = (int)ch(dec[t0], c), b = (int)ch(dec[t0 + 1], c), d = (int)ch(dec[t0 + 4], c), e = (int)ch(dec[t0 + 5], c);
      qs[q][c] = a + b + d + e;
      const int m0 = a < b ? a : b, m1 = d < e ? d : e, x0 = a > b ? a : b, x1 = d > e ? d : e;
      qmn[q][c] = m0 < m1 ? m0 : m1; qmx[q][c] = x0 > x1 ? x0 : x1;
    }
  }
  uint32_t cost = 0;
#pragma unroll
  for (uint32_t v = 0; v < 4; v++) {
    const uint32_t qa = (v == 0) ? 0u : (v == 1) ? 1u : (v == 2) ? 0u : 2u, qb = (v == 0) ? 2u : (v == 1) ? 3u : (v == 2) ? 1u : 3u;
    int range = 0, amin = 255, amax = 0;
#pragma unroll
    for (uint32_t c = 0; c < 3; c++) {
      const int avg = (qs[qa][c] + qs[qb][c] + 4) >> 3;
      const int mn = qmn[qa][c] < qmn[qb][c] ? qmn[qa][c] : qmn[qb][c], mx = qmx[qa][c] > qmx[qb][c] ? qmx[qa][c] : qmx[qb][c];
      const int r = (mx - avg) > (avg - mn) ? (mx - avg) : (avg - mn);
      range = r > range ? r : range;
      amin = avg < amin ? avg : amin; amax = avg > amax ? avg : amax;
    }
    const int room = amin < 255 - amax ? amin : 255 - amax;   // a table clamps when its larger modifier exceeds this
    // larger modifiers of the tables: 8, 17, 29, 42, 60, 80, 106, 183
    const uint32_t plain = (room >= 8) + (room >= 17) + (room >= 29) + (room >= 42) + (room >= 60) + (room >= 80) + (room >= 106) + (room >= 183);
    const uint32_t limit = (range > 51) ? 8u : (range >= 7 ? 4u : 2u);
    const uint32_t n_plain = plain < limit ? plain : limit;
    cost += n_plain + 8u * (limit - n_plain);
  }
  // cost: 8 .. 256, halving classes
  uint32_t cls = cost >= 192u ? 0u : cost >= 128u ? 1u : cost >= 96u ? 2u : cost >= 64u ? 3u : cost >= 48u ? 4u : cost >= 32u ? 5u :
                 cost >= 24u ? 6u : cost >= 16u ? 7u : cost >= 12u ? 8u : cost >= 9u ? 9u : 10u;
  return UBT(uastc_mode_has_etc1_bias)[mode] ? cls : 11u;
}

// `lanes` (device): mask of the warp's lanes that make this call together (0: unknown, see etc1_hints_search).
UB_BIG void etc1_hints_job(uint32_t mode, const uint32_t* px, const uint32_t* dec, uint32_t level, uint32_t flags, uint32_t fslot, uint32_t individ,
                          Etc1JobOut& o, uint32_t lanes = 0) {
  const Etc1Plan pl = etc1_plan(mode, level, flags);
  uint32_t first_flip, last_flip;
  etc1_hints_range(pl, flags, dec, first_flip, last_flip);
  o.err = ~0ull; o.hint = 0; o.pad_ = 0;
  const uint32_t flip = first_flip + fslot;
  const bool none = flip >= last_flip || (individ != ETC1_JOB_BOTH_MODES && individ >= pl.last_individ);
#ifdef __CUDA_ARCH__
  if (lanes) lanes = __ballot_sync(lanes, !none);
#endif
  if (none) return;
  Etc1Sub sub[2];
  etc1_hints_setup(sub, flip, px, dec);
  uint32_t f = 0, d = 0, i0 = 0, i1 = 0, bias = 0;
  uint64_t best = ~0ull;
  if (individ == ETC1_JOB_BOTH_MODES) {
#pragma unroll 1
    for (uint32_t i = 0; i < pl.last_individ; i++) etc1_hints_search(pl, sub, flip, i, level, best, f, d, i0, i1, bias, lanes);
  } else etc1_hints_search(pl, sub, flip, individ, level, best, f, d, i0, i1, bias, lanes);
  o.err = best;
  o.hint = f | (d << 1) | (i0 << 2) | (i1 << 5) | (bias << 8);
}
UB_FN void etc1_hints_combine(const Etc1JobOut* jobs, uint32_t n, uint32_t& o_flip, uint32_t& o_diff, uint32_t& o_i0, uint32_t& o_i1, uint32_t& o_bias) {
  uint64_t best = ~0ull;
  uint32_t h = 0;
  for (uint32_t i = 0; i < n; i++) if (jobs[i].err < best) { best = jobs[i].err; h = jobs[i].hint; }
  o_flip = h & 1u; o_diff = (h >> 1) & 1u; o_i0 = (h >> 2) & 7u; o_i1 = (h >> 5) & 7u; o_bias = h >> 8;
}

// ---------------------------------------------------------------------------------------------------------
// ETC2 EAC A8 table / multiplier hint
UB_BIG uint32_t compute_eac_hint(const uint32_t* dec, uint32_t mul_rad, uint32_t table_mask) {
  int a[16];
  int mn = 255, mx = 0;
  for (uint32_t i = 0; i < 16; i++) { a[i] = (int)(dec[i] >> 24); if (a[i] < mn) mn = a[i]; if (a[i] > mx) mx = a[i]; }
  if (mn == mx) return 13u | (1u << 4);
  const uint32_t arange = (uint32_t)(mx - mn);
  uint64_t best = ~0ull;
  uint32_t b_table = 0, b_mul = 0;
  for (uint32_t table = 0; table < 16; table++) {
    if (!((table_mask >> table) & 1u)) continue;
    const int8_t* T = UBT(etc2_eac) + table * 8;
    const float range = (float)((int)T[7] - (int)T[3]);
    const float t = f_div((float)(0 - (int)T[3]), range);
    const int center = (int)roundf(f_add((float)mn, f_mul(f_sub((float)mx, (float)mn), t)));
    const int base = clamp255(center);
    const int mul = (int)roundf(f_div((float)arange, range));
    const int mul_low = clampi(mul - (int)mul_rad, 1, 15), mul_high = clampi(mul + (int)mul_rad, 1, 15);
    for (int m = mul_low; m <= mul_high; m++) {
      uint64_t total = 0;
      for (uint32_t i = 0; i < 16; i++) {
        uint32_t be = 0xFFFFFFFFu;
        for (uint32_t s = 0; s < 8; s++) {
          int e = a[i] - clamp255(m * (int)T[s] + base);
          if (e < 0) e = -e;
          if ((uint32_t)e < be) be = (uint32_t)e;
        }
        total += (uint64_t)be * be;
      }
      if (total < best) { best = total; b_mul = (uint32_t)m; b_table = table; if (!best) return b_table | (b_mul << 4); }
    }
  }
  return b_table | (b_mul << 4);
}

// ---------------------------------------------------------------------------------------------------------
// BC1 hints
struct Bc1Block { uint32_t lo16, hi16, sels; };  // colour0, colour1, 2-bit selectors (texel 0 in the low bits)

UB_FN void bc1_find_sels(const uint32_t* px, uint32_t lr, uint32_t lg, uint32_t lb, uint32_t hr, uint32_t hg, uint32_t hb, uint8_t* sels) {
  int br[4], bg[4], bb[4];
  br[0] = (int)((lr << 3) | (lr >> 2)); bg[0] = (int)((lg << 2) | (lg >> 4)); bb[0] = (int)((lb << 3) | (lb >> 2));
  br[3] = (int)((hr << 3) | (hr >> 2)); bg[3] = (int)((hg << 2) | (hg >> 4)); bb[3] = (int)((hb << 3) | (hb >> 2));
  br[1] = (br[0] * 2 + br[3]) / 3; bg[1] = (bg[0] * 2 + bg[3]) / 3; bb[1] = (bb[0] * 2 + bb[3]) / 3;
  br[2] = (br[3] * 2 + br[0]) / 3; bg[2] = (bg[3] * 2 + bg[0]) / 3; bb[2] = (bb[3] * 2 + bb[0]) / 3;
  int ar = br[3] - br[0], ag = bg[3] - bg[0], ab = bb[3] - bb[0];
  int dots[4];
  for (int i = 0; i < 4; i++) dots[i] = br[i] * ar + bg[i] * ag + bb[i] * ab;
  const int t0 = dots[0] + dots[1], t1 = dots[1] + dots[2], t2 = dots[2] + dots[3];
  ar *= 2; ag *= 2; ab *= 2;
#pragma unroll 4
  for (uint32_t i = 0; i < 16; i++) {
    const int d = (int)ch(px[i], 0) * ar + (int)ch(px[i], 1) * ag + (int)ch(px[i], 2) * ab;
    sels[i] = (uint8_t)(3 - ((d <= t0) + (d < t1) + (d < t2)));
  }
}

UB_FN bool bc1_least_squares(const uint32_t* px, const uint8_t* sels, float* xl, float* xh) {
  uint32_t q00[3] = {0, 0, 0}, tt[3] = {0, 0, 0}, wacc = 0;
#pragma unroll 4
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t s = sels[i];
    wacc += (s == 0) ? 0x000009u : (s == 1) ? 0x010204u : (s == 2) ? 0x040201u : 0x090000u;
    for (uint32_t c = 0; c < 3; c++) { const uint32_t v = ch(px[i], c); tt[c] += v; q00[c] += s * v; }
  }
  const float z00 = (float)((wacc >> 16) & 0xFFu), z10 = (float)((wacc >> 8) & 0xFFu), z11 = (float)(wacc & 0xFFu);
  const float z01 = z10;
  float det = f_sub(f_mul(z00, z11), f_mul(z01, z10));
  if ((det < 0 ? -det : det) < 1e-8f) return false;
  det = f_div(3.0f, det);
  const float iz00 = f_mul(z11, det), iz01 = f_mul(-z01, det), iz10 = f_mul(-z10, det), iz11 = f_mul(z00, det);
  for (uint32_t c = 0; c < 3; c++) {
    const float fq00 = (float)q00[c];
    const float q10 = f_sub(f_mul((float)tt[c], 3.0f), fq00);
    xl[c] = f_add(f_mul(iz00, fq00), f_mul(iz01, q10));
    xh[c] = f_add(f_mul(iz10, fq00), f_mul(iz11, q10));
  }
  for (uint32_t c = 0; c < 3; c++)
    if (xl[c] < 0.0f || xh[c] > 255.0f) {
      uint32_t lo = 0xFFFFFFFFu, hi = 0;
#pragma unroll 1
      for (uint32_t i = 0; i < 16; i++) { lo = minu(lo, ch(px[i], c)); hi = maxu(hi, ch(px[i], c)); }
      if (lo == hi) { xl[c] = (float)lo; xh[c] = (float)hi; }
    }
  return true;
}

UB_FN uint32_t bc1_to5(uint32_t v) { v = v * 31 + 128; return ((v + (v >> 8)) >> 8) & 255u; }
UB_FN uint32_t bc1_to6(uint32_t v) { v = v * 63 + 128; return ((v + (v >> 8)) >> 8) & 255u; }

UB_FN Bc1Block bc1_solid(uint32_t fr, uint32_t fg, uint32_t fb) {
  uint32_t mask = 0xAA;
  const uint32_t m5r = UBT(bc1_match5)[fr], m6g = UBT(bc1_match6)[fg], m5b = UBT(bc1_match5)[fb];
  uint32_t max16 = ((m5r & 255u) << 11) | ((m6g & 255u) << 5) | (m5b & 255u);
  uint32_t min16 = ((m5r >> 8) << 11) | ((m6g >> 8) << 5) | (m5b >> 8);
  if (min16 == max16) {
    mask = 0;
    if (min16 > 0) min16--; else { max16 = 1; min16 = 0; mask = 0x55; }
  }
  if (max16 < min16) { const uint32_t t = max16; max16 = min16; min16 = t; mask ^= 0x55; }
  Bc1Block b; b.lo16 = max16; b.hi16 = min16; b.sels = mask * 0x01010101u;
  return b;
}

// (The 16-texel loops are unrolled by 4, not fully: fully unrolled the function is 34 KB of SASS, more than the 32 KB
// instruction cache, which the free-running warps of the BC1 / BC3 transcode kernels then thrash.)
// encode_bc1 with flags 0 (have_sels = false) or cEncodeBC1UseSelectors (sels given, linear 0..3); ls_passes = 1, or 2 for
// cEncodeBC1HighQuality (XC:17807)
UB_BIG Bc1Block encode_bc1(const uint32_t* px, bool have_sels, uint8_t* sels, uint32_t ls_passes = 1) {
  int avg_r = -1, avg_g = 0, avg_b = 0;
  uint32_t lr = 0, lg = 0, lb = 0, hr = 0, hg = 0, hb = 0;
  if (!have_sels) {
    const uint32_t f0 = px[0] & 0xFFFFFFu;
    bool solid = true;
    for (uint32_t j = 1; j < 16; j++) if ((px[j] & 0xFFFFFFu) != f0) { solid = false; break; }
    if (solid) return bc1_solid(ch(px[0], 0), ch(px[0], 1), ch(px[0], 2));
    int tot[3] = {0, 0, 0}, mx[3] = {0, 0, 0}, mn[3] = {255, 255, 255};
#pragma unroll 4
    for (uint32_t i = 0; i < 16; i++)
      for (uint32_t c = 0; c < 3; c++) { const int v = (int)ch(px[i], c); tot[c] += v; if (v > mx[c]) mx[c] = v; if (v < mn[c]) mn[c] = v; }
    avg_r = (tot[0] + 8) >> 4; avg_g = (tot[1] + 8) >> 4; avg_b = (tot[2] + 8) >> 4;
    int icov[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll 4
    for (uint32_t i = 0; i < 16; i++) {
      const int r = (int)ch(px[i], 0) - avg_r, g = (int)ch(px[i], 1) - avg_g, b = (int)ch(px[i], 2) - avg_b;
      icov[0] += r * r; icov[1] += r * g; icov[2] += r * b; icov[3] += g * g; icov[4] += g * b; icov[5] += b * b;
    }
    float cov[6];
    const float k255 = f_div(1.0f, 255.0f);
    for (int i = 0; i < 6; i++) cov[i] = f_mul((float)icov[i], k255);
    float xr = (float)(mx[0] - mn[0]), xg = (float)(mx[1] - mn[1]), xb = (float)(mx[2] - mn[2]);
    for (int it = 0; it < 4; it++) {
      const float r = f_add(f_add(f_mul(xr, cov[0]), f_mul(xg, cov[1])), f_mul(xb, cov[2]));
      const float g = f_add(f_add(f_mul(xr, cov[1]), f_mul(xg, cov[3])), f_mul(xb, cov[4]));
      const float b = f_add(f_add(f_mul(xr, cov[2]), f_mul(xg, cov[4])), f_mul(xb, cov[5]));
      xr = r; xg = g; xb = b;
    }
    const float ar = xr < 0 ? -xr : xr, ag = xg < 0 ? -xg : xg, ab = xb < 0 ? -xb : xb;
    float k = ar > ag ? ar : ag;
    k = k > ab ? k : ab;
    int sr = 306, sg = 601, sb = 117;
    if (k >= 2) { const float m = f_div(1024.0f, k); sr = (int)f_mul(xr, m); sg = (int)f_mul(xg, m); sb = (int)f_mul(xb, m); }
    int low_dot = 0x7FFFFFFF, high_dot = (int)0x80000000, low_c = 0, high_c = 0;
#pragma unroll 4
    for (uint32_t i = 0; i < 16; i++) {
      const int dot = (int)ch(px[i], 0) * sr + (int)ch(px[i], 1) * sg + (int)ch(px[i], 2) * sb;
      if (dot < low_dot) { low_dot = dot; low_c = (int)i; }
      if (dot > high_dot) { high_dot = dot; high_c = (int)i; }
    }
    lr = bc1_to5(ch(px[low_c], 0)); lg = bc1_to6(ch(px[low_c], 1)); lb = bc1_to5(ch(px[low_c], 2));
    hr = bc1_to5(ch(px[high_c], 0)); hg = bc1_to6(ch(px[high_c], 1)); hb = bc1_to5(ch(px[high_c], 2));
    bc1_find_sels(px, lr, lg, lb, hr, hg, hb, sels);
  }
  for (uint32_t ls_pass = 0; ls_pass < ls_passes; ls_pass++) {
    float xl[3], xh[3];
    if (!bc1_least_squares(px, sels, xl, xh)) {
      if (avg_r < 0) {
        int tot[3] = {0, 0, 0};
#pragma unroll 1
        for (uint32_t i = 0; i < 16; i++) for (uint32_t c = 0; c < 3; c++) tot[c] += (int)ch(px[i], c);
        avg_r = (tot[0] + 8) >> 4; avg_g = (tot[1] + 8) >> 4; avg_b = (tot[2] + 8) >> 4;
      }
      lr = UBT(bc1_match5)[avg_r] & 255u; lg = UBT(bc1_match6)[avg_g] & 255u; lb = UBT(bc1_match5)[avg_b] & 255u;
      hr = UBT(bc1_match5)[avg_r] >> 8; hg = UBT(bc1_match6)[avg_g] >> 8; hb = UBT(bc1_match5)[avg_b] >> 8;
    } else {
      const float k5 = f_div(31.0f, 255.0f), k6 = f_div(63.0f, 255.0f);
      lr = (uint32_t)clampi((int)f_add(f_mul(xl[0], k5), .5f), 0, 31); lg = (uint32_t)clampi((int)f_add(f_mul(xl[1], k6), .5f), 0, 63);
      lb = (uint32_t)clampi((int)f_add(f_mul(xl[2], k5), .5f), 0, 31);
      hr = (uint32_t)clampi((int)f_add(f_mul(xh[0], k5), .5f), 0, 31); hg = (uint32_t)clampi((int)f_add(f_mul(xh[1], k6), .5f), 0, 63);
      hb = (uint32_t)clampi((int)f_add(f_mul(xh[2], k5), .5f), 0, 31);
    }
    bc1_find_sels(px, lr, lg, lb, hr, hg, hb, sels);
  }
  uint32_t lc16 = lb | (lg << 5) | (lr << 11), hc16 = hb | (hg << 5) | (hr << 11);
  Bc1Block b;
  if (lc16 == hc16) {
    uint32_t mask = 0;
    if (hc16 > 0) hc16--; else { hc16 = 0; lc16 = 1; mask = 0x55; }
    b.lo16 = lc16; b.hi16 = hc16; b.sels = mask * 0x01010101u;
  } else {
    uint32_t inv = 0;
    if (lc16 < hc16) { const uint32_t t = lc16; lc16 = hc16; hc16 = t; inv = 0x55555555u; }
    uint32_t packed = 0;
#pragma unroll 4
    for (uint32_t i = 0; i < 16; i++) packed |= ((0x78u >> (2 * sels[i])) & 3u) << (i * 2);  // {0,2,3,1}
    b.lo16 = lc16; b.hi16 = hc16; b.sels = packed ^ inv;
  }
  return b;
}

UB_FN void bc1_decode(const Bc1Block& b, uint32_t* out) {  // unpack_bc1, RGB only
  uint32_t c[4][3];
  const uint32_t l = b.lo16, h = b.hi16;
  c[0][0] = (l >> 11) & 31u; c[0][1] = (l >> 5) & 63u; c[0][2] = l & 31u;
  c[1][0] = (h >> 11) & 31u; c[1][1] = (h >> 5) & 63u; c[1][2] = h & 31u;
  for (int k = 0; k < 2; k++) { c[k][0] = (c[k][0] << 3) | (c[k][0] >> 2); c[k][1] = (c[k][1] << 2) | (c[k][1] >> 4); c[k][2] = (c[k][2] << 3) | (c[k][2] >> 2); }
  for (int j = 0; j < 3; j++) {
    if (l > h) { c[2][j] = (c[0][j] * 2 + c[1][j]) / 3; c[3][j] = (c[1][j] * 2 + c[0][j]) / 3; }
    else { c[2][j] = (c[0][j] + c[1][j]) / 2; c[3][j] = 0; }
  }
  for (uint32_t i = 0; i < 16; i++) {
    const uint32_t s = (b.sels >> (2 * i)) & 3u;
    out[i] = c[s][0] | (c[s][1] << 8) | (c[s][2] << 16);
  }
}

UB_FN uint32_t uastc_to_bc1_sel(uint32_t wb, uint32_t w) {  // s_uastc{1..5}_to_bc1 (XC:17365-17369)
  switch (wb) {
    case 1: return w;
    case 2: return (0x78u >> (2 * w)) & 3u;                                   // 0,2,3,1
    case 3: return (0x5FA0u >> (2 * w)) & 3u;                                 // 0,0,2,2,3,3,1,1
    case 4: return (uint32_t)((0x57FFAA80ull >> (2 * w)) & 3u);               // 0,0,0,2,2,2,2,2,3,3,3,3,3,1,1,1
    default: return (uint32_t)((0x555FFFFFAAAAA000ull >> (2 * w)) & 3u);      // 6x0,10x2,10x3,6x1
  }
}

// compute_bc1_hints.  `nc` = the winning candidate AFTER anchor normalisation (what pack + unpack would yield).
UB_BIG void compute_bc1_hints(const Candidate& nc, const uint32_t* px, const uint32_t* dec, uint32_t& hint0, uint32_t& hint1) {
  hint0 = 0; hint1 = 0;
  const uint32_t mode = nc.mode;
  const bool has0 = UBT(uastc_mode_has_bc1_hint0)[mode] != 0, has1 = UBT(uastc_mode_has_bc1_hint1)[mode] != 0;
  if (!has0 && !has1) return;
  const uint32_t d = UBT(uastc_mode_desc)[mode];
  const uint32_t er = md_endpoint_range(d), comps = md_comps(d), wb = md_weight_bits(d), pshift = md_dual_plane(d);
  uint8_t sels[16];
  uint32_t t_px[16], h0_px[16], h1_px[16];
  bc1_decode(encode_bc1(dec, false, sels), t_px);
  for (uint32_t i = 0; i < 16; i++) { h0_px[i] = 0; h1_px[i] = 0; }
  if (has1) {
    // BC1 selector order -> linear order {0,3,1,2}
    for (uint32_t i = 0; i < 16; i++) { const uint32_t s = uastc_to_bc1_sel(wb, nc.w[i << pshift]); sels[i] = (uint8_t)((0x9Cu >> (2 * s)) & 3u); }
    bc1_decode(encode_bc1(dec, true, sels), h1_px);
  }
  if (has0) {
    uint32_t l[3], h[3];
    if (comps == 2) { l[0] = l[1] = l[2] = astc_unquant(er, nc.ep[0]); h[0] = h[1] = h[2] = astc_unquant(er, nc.ep[1]); }
    else for (uint32_t c = 0; c < 3; c++) { l[c] = astc_unquant(er, nc.ep[2 * c]); h[c] = astc_unquant(er, nc.ep[2 * c + 1]); }
    Bc1Block b;
    b.lo16 = ((l[0] * 31 + 127) / 255) << 11 | ((l[1] * 63 + 127) / 255) << 5 | ((l[2] * 31 + 127) / 255);
    b.hi16 = ((h[0] * 31 + 127) / 255) << 11 | ((h[1] * 63 + 127) / 255) << 5 | ((h[2] * 31 + 127) / 255);
    if (b.lo16 == b.hi16) {
      uint32_t mask = 0;
      if (b.hi16 > 0) b.hi16--; else { b.hi16 = 0; b.lo16 = 1; mask = 0x55; }
      b.sels = mask * 0x01010101u;
    } else {
      bool inv = false;
      if (b.lo16 < b.hi16) { const uint32_t t = b.lo16; b.lo16 = b.hi16; b.hi16 = t; inv = true; }
      uint32_t s32 = 0;
      for (int i = 15; i >= 0; --i) { uint32_t s = uastc_to_bc1_sel(wb, nc.w[i << pshift]); if (inv) s ^= 1u; s32 = (s32 << 2) | s; }
      b.sels = s32;
    }
    bc1_decode(b, h0_px);
  }
  uint64_t te = 0, e0 = 0, e1 = 0;
  for (uint32_t i = 0; i < 16; i++) { te += dist_rgb(px[i], t_px[i]); e0 += dist_rgb(px[i], h0_px[i]); e1 += dist_rgb(px[i], h1_px[i]); }
  const float t_err = f_sqrt(u64_to_float_h(te)), t0 = f_sqrt(u64_to_float_h(e0)), t1 = f_sqrt(u64_to_float_h(e1));
  if (has0 && t0 <= f_mul(t_err, 1.075f)) hint0 = 1;
  if (has1 && t1 <= f_mul(t_err, 1.075f)) hint1 = 1;
}

}  // namespace ub
