"""Development probe: time the RDO post-pass on the device and compare with the reference on the same blocks."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import ktx_software_b200 as K
import refutil as R
K.init()
size = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
flags = int(sys.argv[2]) if len(sys.argv) > 2 else (2 | 512)
img = R.img_mix(size, size, 3)
px = R.ref_extract_blocks(img) if R.have_ref() else None
d_px = torch.from_numpy(px).cuda()
n = px.shape[0]
d_u = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
K.encode_blocks_dev(d_px.data_ptr(), d_u.data_ptr(), n, flags)
torch.cuda.synchronize()
u = d_u.cpu().numpy()
for jobs in (4, 0):
    d_w = d_u.clone()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    K.rdo_blocks_dev(d_w.data_ptr(), d_px.data_ptr(), n, flags, rdo_lambda=1.0, total_jobs=jobs)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    got = d_w.cpu().numpy()
    line = "rdo %dx%d jobs=%d: %.1f ms (%.2f us/step/chain, %.1f Mpix/s), modified %d" % (
        size, size, jobs, dt * 1e3, dt * 1e6 / (n / max(1, jobs)), size * size / dt / 1e6, (got != u).any(1).sum())
    if R.have_ref() and n <= 65536:
        t0 = time.perf_counter(); want = R.ref_rdo(u, px, flags, os.cpu_count() or 4, jobs, lam=1.0); rt = time.perf_counter() - t0
        line += "  ref %.1f ms, mismatching blocks %d" % (rt * 1e3, (got != want).any(1).sum())
    print(line, flush=True)
