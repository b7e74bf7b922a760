// uastc_ccc.cuh -- endpoint + weight optimiser for one colour cell (<= 16 texels).
// Restates color_cell_compression and its helpers for the ASTC-range path the UASTC encoder uses (mode == 255,
// perceptual == false, channel weights 1): reference external/basis_universal/encoder/basisu_bc7enc.cpp
//   :1364-1762 color_cell_compression, :1103-1283 find_optimal_solution, :822-1033 evaluate_solution,
//   :394/:460 compute_least_squares_endpoints_rgba/rgb, :1035-1101 degenerate endpoint fix-ups,
//   :625-820 optimal single-colour packers.
// Every float op is spelled through f_*/d_* so rounding matches the reference's -ffp-contract=off SSE build.
#pragma once
#include "ub_common.cuh"

namespace ub {

struct CccParams {
  uint32_t n;          // texels in the cell
  uint32_t nw;         // selector weights: 2,4,8,16,32
  uint32_t range;      // ASTC endpoint range
  bool has_alpha;
  const uint8_t* force_sel;  // RDO endpoint re-fit only; else nullptr
  uint32_t uber_level, ls_passes;
};

const uint32_t CCC_NO_SOLUTION = 0xFFFFFFFFu;  // a cell's error is at most 16 * 4 * 255^2 < 2^23

struct CccResults {
  uint32_t best_err;
  uint32_t lo, hi;            // endpoints as SORTED indices, r | g<<8 | b<<16 | a<<24
  uint32_t astc_lo, astc_hi;  // same endpoints as BISE codes
  alignas(16) uint8_t sel[16];
};

struct Vec4 { float c[4]; };

UB_FN float f_sat(float v) { if (v < 0.0f) v = 0.0f; else if (v > 1.0f) v = 1.0f; return v; }
UB_FN float f_dot4(const Vec4& a, const Vec4& b) {
  return f_add(f_add(f_add(f_mul(a.c[0], b.c[0]), f_mul(a.c[1], b.c[1])), f_mul(a.c[2], b.c[2])), f_mul(a.c[3], b.c[3]));
}
UB_FN void f_normalize4(Vec4& v) {
  float s = f_dot4(v, v);
  if (s != 0.0f) {
    s = f_div(1.0f, f_sqrt(s));
#pragma unroll
    for (int i = 0; i < 4; i++) v.c[i] = f_mul(v.c[i], s);
  }
}

UB_FN const uint8_t* weight_table(uint32_t nw) {
  // g_bc7_weights1 = {0,64} is the two ends of any table
  switch (nw) {
    case 4: return UBT(bc7_weights2);
    case 8: return UBT(bc7_weights3);
    case 16: return UBT(astc_weights4);
    case 32: return UBT(astc_weights5);
    default: return nullptr;
  }
}
UB_FN uint32_t sel_weight(uint32_t nw, uint32_t i) { return (nw == 2) ? (i ? 64u : 0u) : weight_table(nw)[i]; }
UB_FN const uint32_t* lsq_table(uint32_t nw) {
  switch (nw) {
    case 2: return UBT(lsq_w1);
    case 4: return UBT(lsq_w2);
    case 8: return UBT(lsq_w3);
    case 16: return UBT(lsq_w4);
    default: return UBT(lsq_w5);
  }
}

UB_FN uint32_t interp_linear(uint32_t l, uint32_t h, uint32_t w) {  // astc_interpolate_linear (bc7enc.cpp:177)
  l = (l << 8) | l; h = (h << 8) | h;
  return ((l * (64 - w) + h * w + 32) >> 6) >> 8;
}
UB_FN uint32_t sorted_unq(uint32_t range, uint32_t idx) { return UBT(astc_sorted_unquant)[range * 256 + idx] & 255u; }
UB_FN uint32_t sorted_code(uint32_t range, uint32_t idx) { return UBT(astc_sorted_unquant)[range * 256 + idx] >> 8; }
UB_FN uint32_t pk(uint32_t r, uint32_t g, uint32_t b, uint32_t a) { return r | (g << 8) | (b << 16) | (a << 24); }
UB_FN uint32_t ch(uint32_t v, uint32_t c) { return (v >> (8 * c)) & 255u; }

UB_FN uint32_t dist_rgb(uint32_t a, uint32_t b) {
  const int dr = (int)ch(a, 0) - (int)ch(b, 0), dg = (int)ch(a, 1) - (int)ch(b, 1), db = (int)ch(a, 2) - (int)ch(b, 2);
  return (uint32_t)(dr * dr) + (uint32_t)(dg * dg) + (uint32_t)(db * db);
}
UB_FN uint32_t dist_rgba(uint32_t a, uint32_t b) {
  const int da = (int)ch(a, 3) - (int)ch(b, 3);
  return dist_rgb(a, b) + (uint32_t)(da * da);
}

UB_FN void set_astc_endpoints(const CccParams& p, CccResults& r) {
  r.astc_lo = 0; r.astc_hi = 0;
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    r.astc_lo |= sorted_code(p.range, ch(r.lo, c)) << (8 * c);
    r.astc_hi |= sorted_code(p.range, ch(r.hi, c)) << (8 * c);
  }
}

// Sum of squared byte differences of two packed texels, and the dot product of two packed texels (unsigned bytes).
UB_FN uint32_t sqdist4(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  const uint32_t d = __vabsdiffu4(a, b);
  return __dp4a(d, d, 0u);
#else
  uint32_t t = 0;
  for (uint32_t c = 0; c < 4; c++) { const int d = (int)ch(a, c) - (int)ch(b, c); t += (uint32_t)(d * d); }
  return t;
#endif
}
UB_FN uint32_t dot4u(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __dp4a(a, b, 0u);
#else
  return ch(a, 0) * ch(b, 0) + ch(a, 1) * ch(b, 1) + ch(a, 2) * ch(b, 2) + ch(a, 3) * ch(b, 3);
#endif
}

// evaluate_solution (bc7enc.cpp:822): interpolate the palette, project each texel, keep the better of two levels.
// RGB cells run the same code with the alpha byte masked off (the reference's RGB branch ignores alpha throughout).
UB_BIG uint32_t evaluate_solution(uint32_t lo, uint32_t hi, const CccParams& p, const uint32_t* px, CccResults& r) {
  const uint32_t amask = p.has_alpha ? 0xFFFFFFFFu : 0x00FFFFFFu;
  uint32_t amin = 0, amax = 0;
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    amin |= sorted_unq(p.range, ch(lo, c)) << (8 * c);
    amax |= sorted_unq(p.range, ch(hi, c)) << (8 * c);
  }
  amin &= amask; amax &= amask;
  const uint32_t N = p.nw;
  uint32_t wc[32];
  wc[0] = amin;
  wc[N - 1] = amax;
  if (N > 2) {
    // astc_interpolate_linear per channel: ((l*257)*(64-w) + (h*257)*w + 32) >> 14 == (base + delta*w) >> 14
    int base[4], delta[4];
#pragma unroll
    for (uint32_t c = 0; c < 4; c++) {
      const int l = (int)(ch(amin, c) * 257u), h = (int)(ch(amax, c) * 257u);
      base[c] = l * 64 + 32; delta[c] = h - l;
    }
    const uint8_t* wt = weight_table(N);
#pragma unroll 1
    for (uint32_t i = 1; i + 1 < N; i++) {
      const int w = (int)wt[i];
      wc[i] = (uint32_t)((base[0] + delta[0] * w) >> 14) | ((uint32_t)((base[1] + delta[1] * w) >> 14) << 8) |
              ((uint32_t)((base[2] + delta[2] * w) >> 14) << 16) | ((uint32_t)((base[3] + delta[3] * w) >> 14) << 24);
    }
  }

  uint32_t total = 0;
  alignas(16) uint8_t tmp[16];
  if (p.force_sel) {
#pragma unroll 1
    for (uint32_t i = 0; i < p.n; i++) {
      const uint32_t s = p.force_sel[i];
      total += sqdist4(wc[s], px[i] & amask);
      tmp[i] = (uint8_t)s;
    }
  } else {
    // (c - amin) . (amax - amin) == c.amax - c.amin - (amin.amax - amin.amin)
    const int k0 = (int)dot4u(amin, amax) - (int)dot4u(amin, amin);
    const float f = f_div((float)N, f_add((float)(int)sqdist4(amax, amin), .00000125f));
#pragma unroll 1
    for (uint32_t i = 0; i < p.n; i++) {
      const uint32_t c = px[i] & amask;
      const int dot = (int)dot4u(c, amax) - (int)dot4u(c, amin) - k0;
      int s = (int)f_add(f_mul((float)dot, f), .5f);
      s = clampi(s, 1, (int)N - 1);
      const uint32_t e0 = sqdist4(wc[s - 1], c);
      uint32_t e1 = sqdist4(wc[s], c);
      if ((e0 < e1) || (e0 == e1 && s == 1)) { e1 = e0; --s; }
      total += e1;
      tmp[i] = (uint8_t)s;
    }
  }
  if (total < r.best_err) {
    r.best_err = total;
    r.lo = lo; r.hi = hi;
#pragma unroll
    for (uint32_t i = 0; i < 4; i++) ((uint32_t*)r.sel)[i] = ((const uint32_t*)tmp)[i];
  }
  return total;
}

// fixDegenerateEndpoints, mode 255 branch (bc7enc.cpp:1049-1073)
UB_FN void fix_degenerate(uint32_t& lo, uint32_t& hi, const Vec4& xl, const Vec4& xh, uint32_t iscale, int flags) {
#pragma unroll
  for (uint32_t i = 0; i < 3; i++) {
    uint32_t a = ch(lo, i), b = ch(hi, i);
    if (a == b) {
      float d = f_sub(xl.c[i], xh.c[i]);
      if (d < 0) d = -d;
      if (d > 0.000125f) {
        if ((flags & 1) && a > 0) a--;
        if ((flags & 2) && b < iscale) b++;
        lo = (lo & ~(255u << (8 * i))) | (a << (8 * i));
        hi = (hi & ~(255u << (8 * i))) | (b << (8 * i));
      }
    }
  }
}

// find_optimal_solution, ASTC-range branch (bc7enc.cpp:1103-1163), in two halves.  xl/xh in 0..1 units.
// First half: the endpoint pairs it will try, in its order -- one pair, or the four fixDegenerateEndpoints variants.
UB_FN uint32_t solution_pairs(Vec4 xl, Vec4 xh, const CccParams& p, uint32_t* qlo, uint32_t* qhi) {
#pragma unroll
  for (int i = 0; i < 4; i++) { xl.c[i] = f_sat(xl.c[i]); xh.c[i] = f_sat(xh.c[i]); }
  const uint32_t rb = UBT(astc_bise_range)[p.range * 3], rt = UBT(astc_bise_range)[p.range * 3 + 1], rq = UBT(astc_bise_range)[p.range * 3 + 2];
  const uint32_t levels = (1 + 2 * rt + 4 * rq) << rb;
  uint32_t tmin = 0, tmax = 0;
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    const int a = clampi((int)f_add(f_mul(xl.c[c], 255.0f), .5f), 0, 255);
    const int b = clampi((int)f_add(f_mul(xh.c[c], 255.0f), .5f), 0, 255);
    tmin |= (uint32_t)UBT(astc_nearest_sorted_index)[p.range * 256 + a] << (8 * c);
    tmax |= (uint32_t)UBT(astc_nearest_sorted_index)[p.range * 256 + b] << (8 * c);
  }
  bool degenerate = false;
#pragma unroll
  for (uint32_t i = 0; i < 3; i++)
    if (ch(tmin, i) == ch(tmax, i)) {
      float d = f_sub(xl.c[i], xh.c[i]);
      if (d < 0) d = -d;
      if (d > 0.0f) degenerate = true;
    }
  if (!degenerate) { qlo[0] = tmin; qhi[0] = tmax; return 1; }
#pragma unroll
  for (int k = 0; k < 4; k++) {
    uint32_t a = tmin, b = tmax;
    fix_degenerate(a, b, xl, xh, levels - 1, (0x3201 >> (4 * k)) & 15);  // flag order 1, 0, 2, 3
    qlo[k] = a; qhi[k] = b;
  }
  return 4;
}
// Second half: evaluate the pairs in order (a pair equal to the best so far is not evaluated again: it could not improve on
// itself).  ONE call site for every pair of every lane: the lanes of a warp that have an i-th pair evaluate it together, where two
// call sites (one pair here, four there) ran one after the other under partial masks.  4096^2: level 3 60.0 -> 57.7 ms, level 4
// (2048^2) 370 -> 344 ms, level 2 unchanged.  (Listing the pairs of SEVERAL trial refits before evaluating them -- the trials depend
// on each other only through the best-so-far -- was measured too: slower at level 2, no better at 3 and 4; not kept.)
// Returns false once the error is zero.
UB_FN bool evaluate_pairs(const uint32_t* qlo, const uint32_t* qhi, uint32_t n, const CccParams& p, const uint32_t* px, CccResults& r) {
#pragma unroll 1
  for (uint32_t i = 0; i < n; i++) {
    if ((r.best_err == CCC_NO_SOLUTION) || qlo[i] != r.lo || qhi[i] != r.hi) evaluate_solution(qlo[i], qhi[i], p, px, r);
    if (!r.best_err) return false;
  }
  return true;
}
UB_BIG uint32_t find_optimal_solution(Vec4 xl, Vec4 xh, const CccParams& p, const uint32_t* px, CccResults& r) {
  uint32_t qlo[4], qhi[4];
  const uint32_t n = solution_pairs(xl, xh, p, qlo, qhi);
  evaluate_pairs(qlo, qhi, n, p, px, r);
  set_astc_endpoints(p, r);
  return r.best_err;
}

// compute_least_squares_endpoints_rgb / _rgba (bc7enc.cpp:394-518): normal equations.  The reference accumulates
// float products w*c in doubles; w = weight/64 and c <= 255, so every product and every partial sum is an exact
// integer multiple of 1/64 -- integer accumulation followed by one conversion gives the identical double.
UB_BIG void least_squares(const CccParams& p, const uint8_t* sel, const uint32_t* px, Vec4& xl, Vec4& xh) {
  const uint32_t* wt = lsq_table(p.nw);
  const uint8_t* wtab = weight_table(p.nw);   // nullptr for two levels (0, 64)
  const uint32_t nc = p.has_alpha ? 4u : 3u;
  double z00 = 0.0, z10 = 0.0, z11 = 0.0;
  uint32_t qi[4] = {0, 0, 0, 0}, t_rb = 0, t_ga = 0;
#pragma unroll 1
  for (uint32_t i = 0; i < p.n; i++) {
    const uint32_t s = sel[i];
    z00 = d_add(z00, (double)bits_to_float(wt[s * 4 + 0]));
    z10 = d_add(z10, (double)bits_to_float(wt[s * 4 + 1]));
    z11 = d_add(z11, (double)bits_to_float(wt[s * 4 + 2]));
    const uint32_t w = wtab ? (uint32_t)wtab[s] : (s ? 64u : 0u);  // == 64 * wt[s*4+3]
    const uint32_t c = px[i];
#pragma unroll
    for (uint32_t k = 0; k < 4; k++) qi[k] += w * ch(c, k);
    t_rb += c & 0x00FF00FFu;
    t_ga += (c >> 8) & 0x00FF00FFu;
  }
  const uint32_t ti[4] = {t_rb & 0xFFFFu, t_ga & 0xFFFFu, t_rb >> 16, t_ga >> 16};
  const double z01 = z10;
  double det = d_sub(d_mul(z00, z11), d_mul(z01, z10));
  if (det != 0.0) det = d_div(1.0, det);
  const double iz00 = d_mul(z11, det), iz01 = d_mul(-z01, det), iz10 = d_mul(-z10, det), iz11 = d_mul(z00, det);
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    if (c < nc) {
      const double q00 = d_mul((double)qi[c], 0.015625), q10 = d_sub((double)ti[c], q00);
      xl.c[c] = (float)d_add(d_mul(iz00, q00), d_mul(iz01, q10));
      xh.c[c] = (float)d_add(d_mul(iz10, q00), d_mul(iz11, q10));
    }
  }
  if (!p.has_alpha) { xl.c[3] = 255.0f; xh.c[3] = 255.0f; }
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    if (c < nc && ((xl.c[c] < 0.0f) || (xh.c[c] > 255.0f))) {
      uint32_t lo_v = 0xFFFFFFFFu, hi_v = 0;
#pragma unroll 1
      for (uint32_t i = 0; i < p.n; i++) { lo_v = minu(lo_v, ch(px[i], c)); hi_v = maxu(hi_v, ch(px[i], c)); }
      if (lo_v == hi_v) { xl.c[c] = (float)lo_v; xh.c[c] = (float)hi_v; }
    }
  }
}

// Refit with the given selectors and evaluate: the body repeated all over color_cell_compression.
// Returns false when the best error reached zero (the reference returns immediately in that case).
UB_BIG bool refit(const CccParams& p, const uint8_t* sel, const uint32_t* px, CccResults& r) {
  Vec4 xl, xh;
  least_squares(p, sel, px, xl, xh);
  const float k = f_div(1.0f, 255.0f);
#pragma unroll
  for (int c = 0; c < 4; c++) { xl.c[c] = f_mul(xl.c[c], k); xh.c[c] = f_mul(xh.c[c], k); }
  return find_optimal_solution(xl, xh, p, px, r) != 0;
}

// Optimal single-colour packers (bc7enc.cpp:625-820), one routine driven by the (range, weights, alpha) case.
// kind: 0 = (8,8,rgb) 1 = (7,4,rgb) 2 = (8,4,rgba) 3 = (13,4,rgb) 4 = (11,32,rgb); -1 = none for this cell type
UB_FN int single_color_kind(const CccParams& p) {
  if (p.range == 8 && p.nw == 8 && !p.has_alpha) return 0;
  if (p.range == 7 && p.nw == 4 && !p.has_alpha) return 1;
  if (p.range == 8 && p.nw == 4 && p.has_alpha) return 2;
  if (p.range == 13 && p.nw == 4 && !p.has_alpha) return 3;
  if (p.range == 11 && p.nw == 32 && !p.has_alpha) return 4;
  return -1;
}
UB_BIG uint32_t pack_one_color(int kind, const CccParams& p, const uint32_t* px, uint32_t r, uint32_t g, uint32_t b, uint32_t a,
                              uint32_t& lo, uint32_t& hi, uint32_t& opt_sel) {
  const uint32_t* tab; uint32_t fill_a, w;
  switch (kind) {
    case 0: tab = UBT(opt_r8_w3); opt_sel = 2; fill_a = 0; w = UBT(bc7_weights3)[2]; break;
    case 1: tab = UBT(opt_r7_w2); opt_sel = 1; fill_a = 0; w = UBT(bc7_weights2)[1]; break;
    case 2: tab = UBT(opt_r8_w2); opt_sel = 1; fill_a = 0; w = UBT(bc7_weights2)[1]; break;
    case 3: tab = UBT(opt_r13_w2); opt_sel = 1; fill_a = 47; w = UBT(bc7_weights2)[1]; break;
    default: tab = UBT(opt_r11_w5); opt_sel = 13; fill_a = 31; w = UBT(astc_weights5)[13]; break;
  }
  const uint32_t er = tab[r], eg = tab[g], eb = tab[b];
  uint32_t la = fill_a, ha = fill_a;
  if (kind == 2) { const uint32_t ea = tab[a]; la = (ea >> 16) & 255u; ha = ea >> 24; }
  lo = pk((er >> 16) & 255u, (eg >> 16) & 255u, (eb >> 16) & 255u, la);
  hi = pk(er >> 24, eg >> 24, eb >> 24, ha);
  uint32_t pc = 0;
  const uint32_t ncomp = (kind == 0 || kind == 1) ? 3u : 4u;
  for (uint32_t i = 0; i < ncomp; i++)
    pc |= interp_linear(sorted_unq(p.range, ch(lo, i)), sorted_unq(p.range, ch(hi, i)), w) << (8 * i);
  uint32_t total = 0;
  const uint32_t amask = (kind == 2) ? 0xFFFFFFFFu : 0x00FFFFFFu;
#pragma unroll 1
  for (uint32_t i = 0; i < p.n; i++) total += sqdist4(pc & amask, px[i] & amask);
  return total;
}

// color_cell_compression (bc7enc.cpp:1364-1762).  px: p.n texels, r | g<<8 | b<<16 | a<<24.
UB_BIG uint32_t color_cell_compression(const CccParams& p, const uint32_t* px, CccResults& r) {
  r.best_err = CCC_NO_SOLUTION;
  r.lo = 0; r.hi = 0; r.astc_lo = 0; r.astc_hi = 0;
  const int sck = p.force_sel ? -1 : single_color_kind(p);

  // per-channel sums (the reference adds the texels up in floats; the sums are small integers, so this is exact)
  uint32_t t_rb = 0, t_ga = 0, diff = 0;
#pragma unroll 1
  for (uint32_t i = 0; i < p.n; i++) {
    const uint32_t c = px[i];
    t_rb += c & 0x00FF00FFu;
    t_ga += (c >> 8) & 0x00FF00FFu;
    diff |= c ^ px[0];
  }

  if (sck >= 0) {
    const uint32_t mask = (sck == 2) ? 0xFFFFFFFFu : 0x00FFFFFFu;
    if (!(diff & mask)) {
      uint32_t os;
      const uint32_t e = pack_one_color(sck, p, px, ch(px[0], 0), ch(px[0], 1), ch(px[0], 2), ch(px[0], 3), r.lo, r.hi, os);
#pragma unroll
      for (uint32_t i = 0; i < 4; i++) ((uint32_t*)r.sel)[i] = os * 0x01010101u;
      r.best_err = e;
      set_astc_endpoints(p, r);
      return e;
    }
  }

  // mean colour
  Vec4 mean = {{(float)(t_rb & 0xFFFFu), (float)(t_ga & 0xFFFFu), (float)(t_rb >> 16), (float)(t_ga >> 16)}};
  Vec4 mean_scaled;
  {
    const float k0 = f_div(1.0f, (float)p.n), k1 = f_div(1.0f, f_mul((float)p.n, 255.0f));
#pragma unroll
    for (uint32_t c = 0; c < 4; c++) { mean_scaled.c[c] = f_mul(mean.c[c], k0); mean.c[c] = f_sat(f_mul(mean.c[c], k1)); }
  }

  Vec4 axis = {{0, 0, 0, 0}};
  if (p.has_alpha) {
#pragma unroll 1
    for (uint32_t i = 0; i < p.n; i++) {
      Vec4 col;
#pragma unroll
      for (uint32_t c = 0; c < 4; c++) col.c[c] = f_sub((float)ch(px[i], c), mean_scaled.c[c]);
      Vec4 n = i ? axis : col;
      f_normalize4(n);
#pragma unroll
      for (uint32_t k = 0; k < 4; k++) {
        Vec4 t;
#pragma unroll
        for (uint32_t c = 0; c < 4; c++) t.c[c] = f_mul(col.c[c], col.c[k]);
        axis.c[k] = f_add(axis.c[k], f_dot4(t, n));
      }
    }
    f_normalize4(axis);
  } else {
    float cov[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll 1
    for (uint32_t i = 0; i < p.n; i++) {
      const float cr = f_sub((float)ch(px[i], 0), mean_scaled.c[0]);
      const float cg = f_sub((float)ch(px[i], 1), mean_scaled.c[1]);
      const float cb = f_sub((float)ch(px[i], 2), mean_scaled.c[2]);
      cov[0] = f_add(cov[0], f_mul(cr, cr)); cov[1] = f_add(cov[1], f_mul(cr, cg)); cov[2] = f_add(cov[2], f_mul(cr, cb));
      cov[3] = f_add(cov[3], f_mul(cg, cg)); cov[4] = f_add(cov[4], f_mul(cg, cb)); cov[5] = f_add(cov[5], f_mul(cb, cb));
    }
    float xr = .9f, xg = 1.0f, xb = .7f;
#pragma unroll 1
    for (uint32_t it = 0; it < 3; it++) {
      float rr = f_add(f_add(f_mul(xr, cov[0]), f_mul(xg, cov[1])), f_mul(xb, cov[2]));
      float gg = f_add(f_add(f_mul(xr, cov[1]), f_mul(xg, cov[3])), f_mul(xb, cov[4]));
      float bb = f_add(f_add(f_mul(xr, cov[2]), f_mul(xg, cov[4])), f_mul(xb, cov[5]));
      const float ar = rr < 0 ? -rr : rr, ag = gg < 0 ? -gg : gg, ab = bb < 0 ? -bb : bb;
      float m = ar > ag ? ar : ag;
      m = m > ab ? m : ab;
      if (m > 1e-10f) { m = f_div(1.0f, m); rr = f_mul(rr, m); gg = f_mul(gg, m); bb = f_mul(bb, m); }
      xr = rr; xg = gg; xb = bb;
    }
    float len = f_add(f_add(f_mul(xr, xr), f_mul(xg, xg)), f_mul(xb, xb));
    if (!(len < 1e-10f)) {
      len = f_div(1.0f, f_sqrt(len));
      axis.c[0] = f_mul(xr, len); axis.c[1] = f_mul(xg, len); axis.c[2] = f_mul(xb, len); axis.c[3] = 0;
    }
  }
  if (f_dot4(axis, axis) < .5f) {
    axis.c[0] = 1.0f; axis.c[1] = 1.0f; axis.c[2] = 1.0f; axis.c[3] = p.has_alpha ? 1.0f : 0.0f;
    f_normalize4(axis);
  }

  float l = 1e+9f, h = -1e+9f;
#pragma unroll 1
  for (uint32_t i = 0; i < p.n; i++) {
    Vec4 q;
#pragma unroll
    for (uint32_t c = 0; c < 4; c++) q.c[c] = f_sub((float)ch(px[i], c), mean_scaled.c[c]);
    const float d = f_dot4(q, axis);
    l = (l < d) ? l : d;
    h = (h > d) ? h : d;
  }
  const float inv255 = f_div(1.0f, 255.0f);
  l = f_mul(l, inv255); h = f_mul(h, inv255);
  Vec4 minc, maxc;
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    minc.c[c] = f_sat(f_add(mean.c[c], f_mul(axis.c[c], l)));
    maxc.c[c] = f_sat(f_add(mean.c[c], f_mul(axis.c[c], h)));
  }
  {
    const float s0 = f_add(f_add(f_add(minc.c[0], minc.c[1]), minc.c[2]), minc.c[3]);
    const float s1 = f_add(f_add(f_add(maxc.c[0], maxc.c[1]), maxc.c[2]), maxc.c[3]);
    if (s0 > s1) { const Vec4 t = minc; minc = maxc; maxc = t; }
  }

  if (!find_optimal_solution(minc, maxc, p, px, r)) return 0;
#pragma unroll 1
  for (uint32_t i = 0; i < p.ls_passes; i++)
    if (!refit(p, r.sel, px, r)) return 0;

  if (!p.force_sel && p.uber_level > 0) {
    alignas(16) uint8_t st[16], s1[16];
    uint32_t min_sel = 256, max_sel = 0;
#pragma unroll 1
    for (uint32_t i = 0; i < p.n; i++) { st[i] = r.sel[i]; min_sel = minu(min_sel, st[i]); max_sel = maxu(max_sel, st[i]); }
    const int max_selector = (int)p.nw - 1;
    // three perturbations: raise the lowest selector, lower the highest, both (bc7enc.cpp:1573-1641)
#pragma unroll 1
    for (uint32_t trial = 1; trial <= 3; trial++) {
#pragma unroll 1
      for (uint32_t i = 0; i < p.n; i++) {
        uint32_t s = st[i];
        if ((trial & 1u) && s == min_sel && s < p.nw - 1) s++;
        else if ((trial & 2u) && s == max_sel && s > 0) s--;
        s1[i] = (uint8_t)s;
      }
      if (!refit(p, s1, px, r)) return 0;
    }

    const uint32_t thresh = (p.n * 56) >> 4;
    if (p.uber_level >= 2 && r.best_err > thresh) {
      const int Q = (p.uber_level >= 4) ? (int)p.uber_level - 2 : 1;
#pragma unroll 1
      for (int ly = -Q; ly <= 1; ly++)
#pragma unroll 1
        for (int hy = max_selector - 1; hy <= max_selector + Q; hy++) {
          if (ly == 0 && hy == max_selector) continue;
          const float den = f_sub((float)hy, (float)ly);
#pragma unroll 1
          for (uint32_t i = 0; i < p.n; i++) {
            const float num = f_mul((float)max_selector, f_sub((float)st[i], (float)ly));
            float v = floorf(f_add(f_div(num, den), .5f));
            if (v < 0) v = 0; else if (v > (float)max_selector) v = (float)max_selector;
            s1[i] = (uint8_t)v;
          }
          if (!refit(p, s1, px, r)) return 0;
        }
    }
  }

  if (sck >= 0) {
    const uint32_t mr = (uint32_t)(int)f_add(.5f, f_mul(mean.c[0], 255.0f)), mg = (uint32_t)(int)f_add(.5f, f_mul(mean.c[1], 255.0f));
    const uint32_t mb = (uint32_t)(int)f_add(.5f, f_mul(mean.c[2], 255.0f)), ma = (uint32_t)(int)f_add(.5f, f_mul(mean.c[3], 255.0f));
    uint32_t lo, hi, os;
    const uint32_t e = pack_one_color(sck, p, px, mr, mg, mb, ma, lo, hi, os);
    if (e < r.best_err) {
      r.best_err = e; r.lo = lo; r.hi = hi;
#pragma unroll
      for (uint32_t i = 0; i < 4; i++) ((uint32_t*)r.sel)[i] = os * 0x01010101u;
      set_astc_endpoints(p, r);
    }
  }
  return r.best_err;
}

}  // namespace ub
