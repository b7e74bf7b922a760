// uastc_hints.cuh -- transcoding hints stored in a UASTC block (BC1 hints, ETC1 flip/diff/intensity/bias, ETC2 EAC
// table+multiplier).  Restates compute_bc1_hints, compute_etc1_hints, uastc_pack_eac_a8 and the solid-colour ETC1
// packer: reference external/basis_universal/encoder/basisu_uastc_enc.cpp:2536-3105, encoder/basisu_etc.cpp:181.
#pragma once
#include "uastc_modes.cuh"

namespace ub {

struct BlockHints;
struct EncCfg;

#ifndef UB_HINTS_IMPLEMENTED
// Stage-A placeholders: colour data is complete, hint fields are zero (tests mask the hint bits until stage B lands).
UB_FN void hints_solid_etc1(uint32_t rgba, uint32_t& diff, uint32_t& inten, uint32_t& sel, uint32_t& r, uint32_t& g, uint32_t& b) {
  (void)rgba; diff = 0; inten = 0; sel = 0; r = 0; g = 0; b = 0;
}
#define UB_COMPUTE_HINTS_STUB 1
#endif

}  // namespace ub
