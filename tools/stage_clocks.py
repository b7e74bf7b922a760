"""Development aid: per-stage busy / wall cycles of k_encode_jobs from a -DEJ_STAGE_CLOCKS=1 build
(tools/build_variant.sh clk -DEJ_STAGE_CLOCKS=1; B200_UASTC_LIB=.../libuastc_clk.so python tools/stage_clocks.py [size] [level])."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import ktx_software_b200 as K
from ktx_software_b200 import api
import refutil as R

K.init()
size = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
lvl = int(sys.argv[2]) if len(sys.argv) > 2 else 2
img = R.img_mix(size, size, 1)
px = torch.from_numpy(img.reshape(size // 4, 4, size // 4, 4, 4).transpose(0, 2, 1, 3, 4).reshape(-1, 64).copy()).cuda()
n = px.shape[0]
out = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
lib = api.lib()
buf = (ctypes.c_ulonglong * 20)()
K.encode_blocks_dev(px.data_ptr(), out.data_ptr(), n, lvl, st); torch.cuda.synchronize()
lib.b200_uastc_debug_stage_clocks(buf, 1)
K.encode_blocks_dev(px.data_ptr(), out.data_ptr(), n, lvl, st); torch.cuda.synchronize()
lib.b200_uastc_debug_stage_clocks(buf, 1)
names = ["claim", "load", "lists", "estimate", "cells", "score", "head+sort", "head end", "etc1 jobs", "tail"]
busy, wall = list(buf[:10]), list(buf[10:])
tw = sum(wall)
print("%dx%d level %d" % (size, size, lvl))
for k in range(10):
    print("%-10s wall %5.1f%%  busy/wall %5.1f%%  idle share of kernel %4.1f%%" % (names[k], 100.0 * wall[k] / tw, 100.0 * busy[k] / max(1, wall[k]), 100.0 * (wall[k] - busy[k]) / tw))
