// port_host.cpp -- TEST INFRASTRUCTURE ONLY.  Host (g++) build of the product's own algorithm headers
// (ktx_software_b200/csrc/*.cuh compile as plain C++ when __CUDACC__ is absent).  It exists so that the CUDA
// restatement can be diffed against the unmodified reference (oracle/_ref/libref_uastc.so) on a box without a
// GPU, and so that bench.py's cpu_baseline can report a "port" figure.  The product library never links this.
#include "../ktx_software_b200/csrc/uastc_to_bc7.cuh"
#include "../ktx_software_b200/csrc/uastc_to_astc.cuh"
#ifdef UB_HAVE_ENCODER
#include "../ktx_software_b200/csrc/uastc_encode.cuh"
#endif

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
