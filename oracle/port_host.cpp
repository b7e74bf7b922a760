// port_host.cpp -- TEST INFRASTRUCTURE ONLY.  Host (g++) build of the product's own algorithm headers
// (ktx_software_b200/csrc/*.cuh compile as plain C++ when __CUDACC__ is absent).  It exists so that the CUDA
// restatement can be diffed against the unmodified reference (oracle/_ref/libref_uastc.so) on a box without a
// GPU, and so that bench.py's cpu_baseline can report a "port" figure.  The product library never links this.
#include "../ktx_software_b200/csrc/uastc_to_bc7.cuh"
#include "../ktx_software_b200/csrc/uastc_to_astc.cuh"
#include "../ktx_software_b200/csrc/uastc_encode.cuh"

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
