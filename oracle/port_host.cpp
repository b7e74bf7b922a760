// port_host.cpp -- TEST INFRASTRUCTURE ONLY.  Host (g++) build of the product's own algorithm headers
// (ktx_software_b200/csrc/*.cuh compile as plain C++ when __CUDACC__ is absent).  It exists so that the CUDA
// restatement can be diffed against the unmodified reference (oracle/_ref/libref_uastc.so) on a box without a
// GPU, and so that bench.py's cpu_baseline can report a "port" figure.  The product library never links this.
#include "../ktx_software_b200/csrc/uastc_to_bc7.cuh"
#include "../ktx_software_b200/csrc/uastc_to_astc.cuh"
#include "../ktx_software_b200/csrc/uastc_encode.cuh"
#include "../ktx_software_b200/csrc/uastc_to_rgba.cuh"
#include "../ktx_software_b200/csrc/uastc_to_bcn.cuh"
#include "../ktx_software_b200/csrc/uastc_jobs.cuh"

#define PORT_API extern "C" __attribute__((visibility("default")))

// fmt: 0 = BC7, 1 = ASTC.  Returns -1 or the index of the first invalid block (all blocks are processed).
PORT_API int64_t port_transcode(const uint8_t* blocks, uint64_t n, uint32_t fmt, uint8_t* out) {
  int64_t bad = -1;
  for (uint64_t i = 0; i < n; i++) {
    ub::Bits128 s, d;
    memcpy(&s, blocks + i * 16, 16);
    const bool ok = fmt ? ub::transcode_block_to_astc(s, d) : ub::transcode_block_to_bc7(s, d);
    memcpy(out + i * 16, &d, 16);
    if (!ok && bad < 0) bad = (int64_t)i;
  }
  return bad;
}

// The same through the mode-specialised code the CUDA kernels run (mode as a template parameter: constant field positions,
// 32-bit block assembly, SIMD selector remaps), dispatched on the block's mode the way k_transcode_tiled's bins are.
template <int TARGET>
static bool known_mode_dispatch(uint32_t mode, const ub::Bits128& s, ub::Bits128& d) {
  switch (mode) {
#define XC_CASE(M) case M: return ub::transcode_known_mode<TARGET, M, false>(s, d);
    XC_CASE(0) XC_CASE(1) XC_CASE(2) XC_CASE(3) XC_CASE(4) XC_CASE(5) XC_CASE(6) XC_CASE(7) XC_CASE(8) XC_CASE(9)
    XC_CASE(10) XC_CASE(11) XC_CASE(12) XC_CASE(13) XC_CASE(14) XC_CASE(15) XC_CASE(16) XC_CASE(17) XC_CASE(18)
#undef XC_CASE
    default: d.lo = 0; d.hi = 0; return false;
  }
}
PORT_API int64_t port_transcode_known(const uint8_t* blocks, uint64_t n, uint32_t fmt, uint8_t* out) {
  int64_t bad = -1;
  for (uint64_t i = 0; i < n; i++) {
    ub::Bits128 s, d;
    memcpy(&s, blocks + i * 16, 16);
    const uint32_t mode = ub::UBT(uastc_huff_modes)[(uint32_t)s.lo & 127u];
    const bool ok = fmt ? known_mode_dispatch<1>(mode, s, d) : known_mode_dispatch<0>(mode, s, d);
    memcpy(out + i * 16, &d, 16);
    if (!ok && bad < 0) bad = (int64_t)i;
  }
  return bad;
}

// UASTC -> 16 RGBA8 texels per block (64 B), the KTX_TTF_RGBA32 decode.  Returns -1 or the first invalid block.
PORT_API int64_t port_decode_rgba(const uint8_t* blocks, uint64_t n, uint8_t* out_px) {
  int64_t bad = -1;
  for (uint64_t i = 0; i < n; i++) {
    ub::Bits128 s;
    memcpy(&s, blocks + i * 16, 16);
    uint32_t px[16];
    const bool ok = ub::decode_block_rgba(s, px);
    memcpy(out_px + i * 64, px, 64);
    if (!ok && bad < 0) bad = (int64_t)i;
  }
  return bad;
}
// The hint-driven targets (kind = ktx_transcode_fmt_e: 0 ETC1_RGB, 1 ETC2_RGBA, 2 BC1_RGB, 3 BC3_RGBA, 4 BC4_R, 5 BC5_RG);
// out holds 8 or 16 bytes per block.  Returns -1 or the index of the first invalid block.
PORT_API int64_t port_transcode_hinted(const uint8_t* blocks, uint64_t n, uint32_t kind, uint32_t high_quality, uint32_t chan0,
                                       uint32_t chan1, uint8_t* out) {
  int64_t bad = -1;
  const uint32_t ob = ub::hinted_target_bytes(kind);
  for (uint64_t i = 0; i < n; i++) {
    ub::Bits128 s, d;
    memcpy(&s, blocks + i * 16, 16);
    const bool ok = ub::transcode_block_hinted(s, kind, high_quality != 0, chan0, chan1, d);
    memcpy(out + i * ob, &d, ob);
    if (!ok && bad < 0) bad = (int64_t)i;
  }
  return bad;
}

// white-box tap: the ETC2 EAC hint byte of every block (table | multiplier << 4), 0xFFFF where the mode carries none or
// the block is invalid.  The reference asserts multiplier >= 1, so tests drop random-payload blocks with multiplier 0.
PORT_API void port_etc2_hints(const uint8_t* blocks, uint64_t n, uint16_t* out) {
  for (uint64_t i = 0; i < n; i++) {
    ub::Bits128 s;
    memcpy(&s, blocks + i * 16, 16);
    ub::HintedForm h;
    out[i] = 0xFFFF;
    if (!ub::hinted_form_t<-1>(s, h) || h.mode == ub::UASTC_MODE_SOLID) continue;
    const ub::HintFields hf = ub::parse_hints(h.mode, h.hints);
    if (hf.has_alpha) out[i] = (uint16_t)hf.etc2_hints;
  }
}

// texels -> 16-bit targets (kind 14 = RGB565, 15 = BGR565, 16 = RGBA4444)
PORT_API void port_pack_texels16(const uint32_t* px, uint64_t n, uint32_t kind, uint16_t* out) {
  for (uint64_t i = 0; i < n; i++) out[i] = (uint16_t)ub::pack_texel16(px[i], kind);
}
PORT_API uint32_t port_weight_unquant(uint32_t wb, uint32_t l) { return ub::weight_unquant(wb, l); }

// n blocks of 16 RGBA texels (64 B each) -> n UASTC blocks.
PORT_API void port_encode_blocks(const uint8_t* px, uint64_t n, uint32_t flags, uint8_t* out) {
  static thread_local ub::Candidate cands[ub::MAX_CANDIDATES];
  static thread_local uint64_t overall[ub::MAX_CANDIDATES];
  static thread_local float uerr[ub::MAX_CANDIDATES];
  for (uint64_t i = 0; i < n; i++) {
    uint32_t p[16];
    memcpy(p, px + i * 64, 64);
    const ub::Bits128 b = ub::encode_uastc_block(p, flags, cands, overall, uerr);
    memcpy(out + i * 16, &b, 16);
  }
}

// The same encode cut into jobs (csrc/uastc_jobs.cuh: what k_encode_jobs computes), one block after the other (level 4: the j4_* enumeration).
PORT_API void port_encode_blocks_jobs(const uint8_t* px, uint64_t n, uint32_t flags, uint8_t* out) {
  for (uint64_t i = 0; i < n; i++) {
    uint32_t p[16];
    memcpy(p, px + i * 64, 64);
    const ub::Bits128 b = ub::encode_uastc_block_jobs(p, flags);
    memcpy(out + i * 16, &b, 16);
  }
}

// white-box tap: one colour cell -> out[0..3] astc lo, [4..7] astc hi, [8..23] selectors; returns the error
PORT_API uint64_t port_ccc(const uint8_t* px, uint32_t n, uint32_t nw, uint32_t range, int has_alpha, uint32_t uber, uint32_t ls, uint8_t* out) {
  ub::CccParams p;
  p.n = n; p.nw = nw; p.range = range; p.has_alpha = has_alpha != 0; p.force_sel = nullptr; p.uber_level = uber; p.ls_passes = ls;
  uint32_t pp[16];
  memcpy(pp, px, n * 4);
  ub::CccResults r;
  memset(&r, 0, sizeof(r));
  const uint64_t e = ub::color_cell_compression(p, pp, r);
  memcpy(out, &r.astc_lo, 4); memcpy(out + 4, &r.astc_hi, 4); memcpy(out + 8, r.sel, 16);
  return e;
}

// ---- RDO: the serial walk, written the plain way over the product's per-block pieces (uastc_rdo.cuh).
#include "../ktx_software_b200/csrc/uastc_rdo.cuh"
#include <vector>

static int g_port_rdo_check = 0;
static uint64_t g_port_rdo_mismatch = 0;
// check = 1: every table-scored trial is also scored the direct way; returns the number of disagreements so far
PORT_API uint64_t port_rdo_set_check(int check) { g_port_rdo_check = check; return g_port_rdo_mismatch; }

static void port_rdo_range(uint8_t* blocks, const uint8_t* px, uint32_t first_index, uint32_t last_index, const ub::RdoParams& p) {
  uint32_t cap = 16;
  while (cap < 2 * (last_index - first_index) + 16) cap <<= 1;
  std::vector<ub::RdoHistEntry> hist(cap);
  memset(hist.data(), 0, cap * sizeof(ub::RdoHistEntry));
  const uint32_t mask = cap - 1;
  for (uint32_t bi = first_index; bi < last_index; bi++) {
    ub::Bits128 blk;
    memcpy(&blk, blocks + (size_t)bi * 16, 16);
    uint32_t pp[16];
    memcpy(pp, px + (size_t)bi * 64, 64);
    ub::RdoCur c;
    ub::RdoBlockInfo info;
    ub::UnpackedUastc un;
    uint32_t table[ub::RDO_TABLE_WORDS];
    if (!ub::rdo_block_info(blk, info, un)) return;
    if (info.mode != ub::UASTC_MODE_SOLID)
      for (uint32_t l = 0; l < (1u << info.wb); l++) ub::rdo_fill_level(pp, un, info, l, table);
    ub::rdo_prepare_fast(pp, blk, info, table, p, c);
    if (g_port_rdo_check) {  // the table path must agree with the direct definition
      ub::RdoCur c2;
      if (!ub::rdo_prepare(pp, blk, p, c2) || c2.ms_err != c.ms_err || c2.skip != c.skip || c2.sel_lo != c.sel_lo) { g_port_rdo_mismatch++; }
    }
    if (c.solid) continue;
    if (c.skip) { ub::rdo_hist_set(hist.data(), mask, c.first, c.sel_lo, bi); continue; }
    float best_t = ub::rdo_keep_cost(c, ub::rdo_hist_find(hist.data(), mask, c.first, c.sel_lo), bi, p);
    ub::Bits128 best = blk;
    bool modified = false;
    const int lo = (int)first_index > (int)bi - (int)p.window ? (int)first_index : (int)bi - (int)p.window;
    for (int pi = (int)bi - 1; pi >= lo; --pi) {
      ub::Bits128 prev;
      memcpy(&prev, blocks + (size_t)pi * 16, 16);
      uint64_t slo, shi;
      ub::rdo_get_sel(prev, c.first, c.total, slo, shi);
      const int64_t h = ub::rdo_hist_find(hist.data(), mask, c.first, slo);
      const int64_t match = h >= 0 ? h : pi;
      if (match > pi) continue;
      if (ub::UBT(uastc_huff_modes)[(uint32_t)prev.lo & 127u] >= ub::UASTC_TOTAL_MODES) return;  // unpack of prev failed
      float t;
      const bool ok = ub::rdo_trial_fast(table, info, slo, shi, c, bi, (uint32_t)match, p, t);
      if (g_port_rdo_check) {
        float t2; ub::Bits128 tr2;
        const bool ok2 = ub::rdo_trial(pp, blk, prev, c, bi, (uint32_t)match, p, t2, tr2);
        if (ok2 != ok || (ok && t2 != t)) g_port_rdo_mismatch++;
      }
      if (!ok) continue;
      if (t < best_t) { best_t = t; best = ub::rdo_splice(blk, prev, c.first, c.total); modified = true; }
    }
    if (modified) {
      if (p.refine && c.mode == 0) best = ub::rdo_refine_mode0(pp, best);
      best = ub::rdo_rehint(pp, best, p.flags);
      memcpy(blocks + (size_t)bi * 16, &best, 16);
    }
    uint64_t slo, shi;
    ub::rdo_get_sel(best, c.first, c.total, slo, shi);
    ub::rdo_hist_set(hist.data(), mask, c.first, slo, bi);
  }
}

PORT_API void port_rdo(uint8_t* blocks, uint32_t n, const uint8_t* px, float lambda, uint32_t dict_size, float max_rms_ratio,
                       float skip_rms, float smooth_scale, float smooth_std_dev, uint32_t flags, uint32_t total_jobs) {
  ub::RdoParams p;
  p.lambda = lambda; p.max_rms_ratio = max_rms_ratio; p.skip_rms = skip_rms; p.smooth_scale = smooth_scale;
  p.smooth_std_dev = smooth_std_dev; p.window = dict_size / 16 ? dict_size / 16 : 1; p.literal_cost = 100; p.flags = flags; p.refine = 1;
  uint32_t per;
  const uint32_t jobs = ub::rdo_job_ranges(n, total_jobs, &per);
  for (uint32_t j = 0; j < jobs; j++) port_rdo_range(blocks, px, j * per, (j + 1) * per < n ? (j + 1) * per : n, p);
}

// white-box tap: decode_bc7_build(bb) must equal decode_bc7_pixels(bc7_pack(bb)) for every transcodable block
PORT_API int64_t port_check_bc7_build_decode(const uint8_t* blocks, uint64_t n) {
  int64_t bad = 0;
  for (uint64_t i = 0; i < n; i++) {
    ub::Bits128 s;
    memcpy(&s, blocks + i * 16, 16);
    ub::UnpackedUastc u;
    if (!ub::unpack_uastc(s, u, false)) continue;
    ub::Bc7Build bb;
    ub::uastc_to_bc7_build(u, bb);
    uint32_t a[16], b[16];
    ub::decode_bc7_build(bb, a);
    const ub::Bits128 packed = ub::bc7_pack(bb);
    if (!ub::decode_bc7_pixels(packed, b) || memcmp(a, b, 64)) bad++;
  }
  return bad;
}
