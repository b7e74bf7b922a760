// uastc_to_astc.cuh -- UASTC -> ASTC 4x4 block (restates transcode_uastc_to_astc, pack_astc_block, astc_pack_bise,
// pack_astc_solid_block; reference external/basis_universal/transcoder/basisu_transcoder.cpp:15649-15667,
// :14663-14795, :14587-14659, :14482-14501, and the blue-contraction undo in unpack_uastc :15328-15371).
#pragma once
#include "uastc_unpack.cuh"
#include "uastc_to_bc7.cuh"

namespace ub {

// BISE-encode `n` endpoint codes of the given range at bit `pos` (values = bits | trit_or_quint << nbits).
template <int KM>
UB_FN void astc_pack_bise(Bits128& dst, const uint8_t* vals, uint32_t n, uint32_t range, uint32_t pos) {
  const uint32_t nb = (KM >= 0) ? bise_c(range, 0) : UBT(astc_bise_range)[range * 3 + 0];
  const uint32_t trits = (KM >= 0) ? bise_c(range, 1) : UBT(astc_bise_range)[range * 3 + 1];
  const uint32_t quints = (KM >= 0) ? bise_c(range, 2) : UBT(astc_bise_range)[range * 3 + 2];
  const uint32_t bm = (1u << nb) - 1u;
  if (trits) {
    // groups of five: m0 T[1:0] m1 T[3:2] m2 T[4] m3 T[6:5] m4 T[7]
#pragma unroll
    for (uint32_t g = 0; g < 4; g++) {
      if (g * 5 < n) {
        uint32_t v[5];
#pragma unroll
        for (uint32_t i = 0; i < 5; i++) v[i] = (g * 5 + i < n) ? vals[(g * 5 + i) < 18 ? (g * 5 + i) : 17] : 0u;
        const uint32_t t = (v[0] >> nb) + (v[1] >> nb) * 3 + (v[2] >> nb) * 9 + (v[3] >> nb) * 27 + (v[4] >> nb) * 81;
        const uint32_t T = UBT(astc_trit_encode)[t];
        uint64_t w = (v[0] & bm) | ((uint64_t)(T & 3u) << nb) | ((uint64_t)(v[1] & bm) << (2 + nb));
        uint32_t p = 2 * nb + 2;
        w |= (uint64_t)((T >> 2) & 3u) << p;
        w |= (uint64_t)(v[2] & bm) << (p + 2);
        w |= (uint64_t)((T >> 4) & 1u) << (p + 2 + nb);
        w |= (uint64_t)(v[3] & bm) << (p + 3 + nb);
        w |= (uint64_t)((T >> 5) & 3u) << (p + 3 + 2 * nb);
        w |= (uint64_t)(v[4] & bm) << (p + 5 + 2 * nb);
        w |= (uint64_t)((T >> 7) & 1u) << (p + 5 + 3 * nb);
        bits_or(dst, pos, w);
        pos += 5 * nb + 8;
      }
    }
  } else if (quints) {
    // groups of three: m0 Q[2:0] m1 Q[4:3] m2 Q[6:5]
#pragma unroll
    for (uint32_t g = 0; g < 6; g++) {
      if (g * 3 < n) {
        uint32_t v[3];
#pragma unroll
        for (uint32_t i = 0; i < 3; i++) v[i] = (g * 3 + i < n) ? vals[(g * 3 + i) < 18 ? (g * 3 + i) : 17] : 0u;
        const uint32_t q = (v[0] >> nb) + (v[1] >> nb) * 5 + (v[2] >> nb) * 25;
        const uint32_t Q = UBT(astc_quint_encode)[q];
        uint64_t w = (v[0] & bm) | ((uint64_t)(Q & 7u) << nb) | ((uint64_t)(v[1] & bm) << (3 + nb)) |
                     ((uint64_t)((Q >> 3) & 3u) << (3 + 2 * nb)) | ((uint64_t)(v[2] & bm) << (5 + 2 * nb)) |
                     ((uint64_t)((Q >> 5) & 3u) << (5 + 3 * nb));
        bits_or(dst, pos, w);
        pos += 3 * nb + 7;
      }
    }
  } else {
#pragma unroll
    for (uint32_t i = 0; i < 18; i++)
      if (i < n) { bits_or(dst, pos, vals[i]); pos += nb; }
  }
}

UB_FN Bits128 astc_pack_solid(uint32_t rgba) {  // void-extent block, UNORM16 = c | c<<8
  Bits128 d;
  d.lo = 0xFFFFFFFFFFFFFDFCull;
  d.hi = 0;
#pragma unroll
  for (uint32_t c = 0; c < 4; c++) {
    const uint64_t v = (rgba >> (8 * c)) & 255u;
    d.hi |= (v | (v << 8)) << (16 * c);
  }
  return d;
}

template <int KM, bool TS = false>
UB_FN bool transcode_block_to_astc_t(const Bits128& src, Bits128& dst) {
  UnpackedUastc u;
  if (!unpack_uastc_t<KM, TS>(src, u, false)) { dst.lo = 0; dst.hi = 0; return false; }
  const uint32_t mode = (KM >= 0) ? (uint32_t)KM : u.mode;
  if (mode == UASTC_MODE_SOLID) { dst = astc_pack_solid(u.solid); return true; }

  const uint32_t d = (KM >= 0) ? mode_desc_c(KM) : u.desc;
  const uint32_t subsets = md_subsets(d), comps = md_comps(d), er = md_endpoint_range(d), wb = md_weight_bits(d);
  const uint32_t planes = 1 + md_dual_plane(d);

  // UASTC stores endpoints without ASTC blue contraction: order each RGB pair so that sum(e0) <= sum(e1), flipping
  // the subset's weights when we swap (XC:15328-15371).  LA modes (comps == 2) are untouched.
  if (comps >= 3) {
    uint32_t inv = 0;
#pragma unroll
    for (uint32_t s = 0; s < 3; s++) {
      if (s < subsets) {
        const uint32_t base = s * comps * 2;
        // subsets*comps*2 <= 18 and the loop is unrolled, so every index below is a compile-time pattern per (s, comps)
        int s0 = 0, s1 = 0;
#pragma unroll
        for (uint32_t c = 0; c < 3; c++) {
          s0 += (int)astc_unquant_s<TS>(er, u.ep[(base + 2 * c) % 18]);
          s1 += (int)astc_unquant_s<TS>(er, u.ep[(base + 2 * c + 1) % 18]);
        }
        if (s1 < s0) {
          inv |= 1u << s;
#pragma unroll
          for (uint32_t c = 0; c < 4; c++)
            if (c < comps) {
              const uint32_t i0 = (base + 2 * c) % 18, i1 = (base + 2 * c + 1) % 18;
              const uint8_t t = u.ep[i0]; u.ep[i0] = u.ep[i1]; u.ep[i1] = t;
            }
        }
      }
    }
    if (inv) {
      const Bits128 m = weight_invert_mask(u, inv);
      u.w.lo ^= m.lo;
      u.w.hi ^= m.hi;
    }
  }

  Bits128 o;
  o.lo = UBT(uastc_mode_astc_block_mode)[mode];
  o.hi = 0;
  uint32_t pos = 11;
  o.lo |= (uint64_t)(subsets - 1) << pos; pos += 2;
  const uint32_t cem = (comps == 2) ? 4u : (comps == 3) ? 8u : 12u;
  if (subsets == 1) { o.lo |= (uint64_t)cem << pos; pos += 4; }
  else {
    uint32_t seed;
    if (subsets == 3) seed = (UBT(common3)[u.pattern] >> 8) & 0xFFFFu;
    else if (mode == 7) seed = (UBT(common7)[u.pattern] >> 8) & 0xFFFFu;
    else seed = (UBT(common2)[u.pattern] >> 8) & 0xFFFFu;
    o.lo |= (uint64_t)seed << pos; pos += 10;
    o.lo |= (uint64_t)((cem << 2) & 63u) << pos; pos += 6;
  }
  const uint32_t total_wbits = 16 * planes * wb;
  if (planes == 2) bits_or(o, 128 - total_wbits - 2, u.ccs);

  astc_pack_bise<KM>(o, u.ep, comps * 2 * subsets, er, pos);

  // weights: each entry bit-reversed, entry 0 ending at bit 127 -> the reversal of the whole packed string
  {
    // u.w occupies bits [0,total_wbits): reverse 128 bits => lands at [128-total_wbits,128)
    const uint64_t rlo = brev64(u.w.hi), rhi = brev64(u.w.lo);
    o.lo |= rlo;
    o.hi |= rhi;
  }
  dst = o;
  return true;
}
UB_FN bool transcode_block_to_astc(const Bits128& src, Bits128& dst) { return transcode_block_to_astc_t<-1>(src, dst); }

// Mode-dispatched transcode of one block whose mode the caller already decoded (blocks binned by mode, so a warp takes
// one case).  TARGET 0 = BC7, 1 = ASTC.
template <int TARGET, int KM, bool TS = false>
UB_FN bool transcode_known_mode(const Bits128& src, Bits128& dst) {
  return TARGET ? transcode_block_to_astc_t<KM, TS>(src, dst) : transcode_block_to_bc7_t<KM, TS>(src, dst);
}

}  // namespace ub
