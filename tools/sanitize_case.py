#!/usr/bin/env python3
"""The smallest run that launches every kernel family of the library once, for compute-sanitizer (memcheck / racecheck):
job-based encode (levels 0-3), block-at-a-time encode (level 4), RDO chain, every transcode target, texel decode, mip chain.
Checks results against the committed goldens so a sanitizer-clean run is also a correct one."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import ktx_software_b200 as K
import refutil as R

K.init()
g = np.load(os.path.join(ROOT, "tests", "golden", "transcode_golden.npz"))
e = np.load(os.path.join(ROOT, "tests", "golden", "encode_golden.npz"))
h = np.load(os.path.join(ROOT, "tests", "golden", "hinted_golden.npz"))
n = int(os.environ.get("SAN_BLOCKS", "4500"))   # > one 4096-block transcode tile, so the double buffer and the tail tile both run
u = g["uastc"][:n]
assert (K.transcode(u, K.TTF_BC7_RGBA).reshape(-1, 16) == g["bc7"][:n]).all()
assert (K.transcode(u, K.TTF_ASTC_4x4_RGBA).reshape(-1, 16) == g["astc"][:n]).all()
uh = g["uastc"][h["index"][:1100]]
for tf, name in ((K.TTF_BC1_RGB, "bc1"), (K.TTF_BC3_RGBA, "bc3"), (K.TTF_BC4_R, "bc4"), (K.TTF_BC5_RG, "bc5"), (K.TTF_ETC1_RGB, "etc1"),
                 (K.TTF_ETC2_RGBA, "etc2"), (K.TTF_ETC2_EAC_R11, "r11"), (K.TTF_ETC2_EAC_RG11, "rg11")):
    if name in h.files:
        got = K.transcode(uh, tf, decode_flags=0).reshape(-1, K.target_block_bytes(tf))
        assert (got == h[name][:1100]).all(), name
K.decode(u[:24 * 16], K.TTF_RGBA32, [(0, 96, 64)])
K.decode(u[:24 * 16], K.TTF_RGB565, [(0, 94, 61)])
px = e["pixels"][:832]
for lvl in (0, 2, 3):
    assert (K.encode_blocks(px, lvl) == e["level%d" % lvl][:832]).all(), lvl
K.encode_blocks(px[:64], 4)
img = R.img_mix(72, 52, 3)
K.encode_images([img, R.img_mix(5, 3, 1)], 1 | 512, rdo=True, rdo_lambda=2.0, total_jobs=4)
K.encode_mipchain(img, 7, 1)
K.shutdown()
print("sanitize_case ok: launches =", K.launch_count())
