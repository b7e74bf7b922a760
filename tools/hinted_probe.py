"""Development aid: device-resident throughput of the hint-driven transcode targets on a C5-like corpus (GB/s counted as
16 B read + 8 or 16 B written per block)."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ktx_software_b200 as K
K.init()
g = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "transcode_golden.npz"))
h = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "hinted_golden.npz"))
u = g["uastc"][:16384]   # reference-encoded part of the corpus
n = 1 << 24
rep = torch.from_numpy(np.tile(u, (n // u.shape[0], 1))).cuda()
want = sys.argv[1:]
for tf, name in [(6, "bc7"), (10, "astc"), (2, "bc1"), (3, "bc3"), (4, "bc4"), (5, "bc5"), (0, "etc1"), (1, "etc2"), (20, "r11"), (21, "rg11")]:
    if want and name not in want:
        continue
    ob = K.target_block_bytes(tf)
    dst = torch.empty(n * ob, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        K.transcode_dev_ex(tf, 0, True, rep.data_ptr(), dst.data_ptr(), n, st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        K.transcode_dev_ex(tf, 0, True, rep.data_ptr(), dst.data_ptr(), n, st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("%-5s %7.3f ms  %7.1f GB/s  %7.1f Gblocks/s" % (name, ms, n * (16 + ob) / ms / 1e6, n / ms / 1e6))
