#!/usr/bin/env python3
"""Development aid: time k_transcode_tiled<2> (UASTC -> RGBA32 texels) on device-resident blocks.
usage: decode_probe.py [edge=8192] [target=13]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import ktx_software_b200 as K
K.init()
edge = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
tf = int(sys.argv[2]) if len(sys.argv) > 2 else 13
g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "transcode_golden.npz"))
n = (edge // 4) ** 2
rng = np.random.default_rng(1)
for name, idx in (("shuffled corpus", rng.integers(0, g["uastc"].shape[0], n)), ("corpus order", np.arange(n) % g["uastc"].shape[0])):
    src = torch.from_numpy(g["uastc"][idx]).cuda()
    bpp = 4 if tf == 13 else 2
    dst = torch.empty(edge * edge * bpp, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        K.decode_image_dev(tf, src.data_ptr(), dst.data_ptr(), edge, edge, 0, st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    reps = 10
    for _ in range(reps):
        K.decode_image_dev(tf, src.data_ptr(), dst.data_ptr(), edge, edge, 0, st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    byts = n * (16 + 16 * bpp)
    print("%s: %dx%d target %d: %.3f ms, %.1f GB/s algorithmic (%d B/block), %.1f Gpix/s" % (name, edge, edge, tf, ms, byts / ms / 1e6, 16 + 16 * bpp, edge * edge / ms / 1e6))
