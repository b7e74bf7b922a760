import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import ktx_software_b200 as K
import refutil as R
K.init()
px = np.concatenate([R.adversarial_blocks(9, 24), R.ref_extract_blocks(R.img_mix(96, 96, 2))])
for flags in (2 | 512, 3 | 512, 1 | 512):
    a = K.encode_blocks(px, flags); b = R.ref_encode_blocks(px, flags); c = R.port_encode_blocks(px, flags)
    bad = np.nonzero((a != b).any(1))[0]
    print("flags", flags, "bad", len(bad), "of", len(px), "port==ref", (b == c).all())
    for k in bad[:6]:
        print(" blk", k, "gpu mode", R.block_modes(a[k:k+1])[0], "ref mode", R.block_modes(b[k:k+1])[0])
        print("  px ", px[k].tobytes().hex()); print("  gpu", a[k].tobytes().hex()); print("  ref", b[k].tobytes().hex())
