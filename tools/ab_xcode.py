"""Development aid: BC7 / ASTC device-resident transcode time of the library named by B200_UASTC_LIB (or the in-tree build),
through b200_uastc_transcode_dev only, so builds from before the _ex entry points can be compared on the same box."""
import ctypes, os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
path = os.environ.get("B200_UASTC_LIB") or os.path.join(os.path.dirname(__file__), "..", "ktx_software_b200", "libuastc_b200.so")
L = ctypes.CDLL(path)
L.b200_uastc_transcode_dev.argtypes = [ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p]
assert L.b200_uastc_init() == 0
g = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "transcode_golden.npz"))
u = g["uastc"][:16384]
n = 1 << 26
rep = torch.from_numpy(np.tile(u, (n // u.shape[0], 1))).cuda()
if "shuffled" in sys.argv[1:]:   # every tile sees the corpus' whole mode mix (as bench.py's C5 input does)
    rep = rep[torch.randperm(n, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))].contiguous()
dst = torch.empty_like(rep)
st = torch.cuda.current_stream().cuda_stream
for tf, name in [(6, "bc7"), (10, "astc")]:
    for _ in range(3):
        L.b200_uastc_transcode_dev(tf, rep.data_ptr(), dst.data_ptr(), n, st, None)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        L.b200_uastc_transcode_dev(tf, rep.data_ptr(), dst.data_ptr(), n, st, None)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(("shuffled " if "shuffled" in sys.argv[1:] else "") + "%s %-5s %7.3f ms  %7.1f GB/s" % (os.path.basename(path), name, ms, n * 32 / ms / 1e6))
