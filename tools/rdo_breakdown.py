"""Development aid: C3-style encode + RDO through the host API, once warm-up and once timed (wall clock); run under
`ncu --metrics gpu__time_duration.sum` for the per-kernel split."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import ktx_software_b200 as K
import refutil as R
size = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
flags = int(sys.argv[2]) if len(sys.argv) > 2 else (2 | 512)
K.init()
img = R.img_mix(size, size, 3)
K.encode_images([R.img_mix(64, 64, 1)], flags, rdo=True, rdo_lambda=1.0, total_jobs=4)
t0 = time.perf_counter()
out, _ = K.encode_images([img], flags, rdo=True, rdo_lambda=1.0, total_jobs=4)
dt = time.perf_counter() - t0
n = out[0].shape[0]
print("RDO %dx%d flags %d: %.1f ms total, %.2f us per chain step (%d steps per chain)" % (size, size, flags, dt * 1e3, dt * 1e6 / (n / 4), n // 4))
