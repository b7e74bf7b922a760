"""Development probe: wall time of the host-buffer transcode call (pinned buffers), per call."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import ktx_software_b200 as K
K.init()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5592407
g = np.load(os.path.join(ROOT, "tests", "golden", "transcode_golden.npz"))["uastc"]
src = torch.empty((n, 16), dtype=torch.uint8).pin_memory()
src.numpy()[:] = np.resize(g, (n, 16))
dst = torch.empty((n, 16), dtype=torch.uint8).pin_memory()
dst.zero_()
for tgt, name in ((K.TTF_BC7_RGBA, "bc7"), (K.TTF_ASTC_4x4_RGBA, "astc"), (K.TTF_BC7_RGBA, "bc7"), (K.TTF_ASTC_4x4_RGBA, "astc")):
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        K.transcode(src.numpy().reshape(-1), tgt, spans=[(0, 0, n)], dst=dst.numpy().reshape(-1))
        ts.append(1e3 * (time.perf_counter() - t0))
    print(name, " ".join("%.2f" % t for t in ts), "ms;  %.1f GB/s (in+out)" % (32 * n / min(ts) / 1e6))
