"""Quick device-side timing of the encode kernel per level (not the bench; a development probe)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import ktx_software_b200 as K
import refutil as R

K.init()
size = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
levels = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 1, 2, 3]
kinds = os.environ.get("PROBE_KINDS", "mix,noise,smooth").split(",")
mk = {"mix": R.img_mix, "noise": R.img_noise, "smooth": R.img_smooth}
for name, img in ((k, mk[k](size, size, 1)) for k in kinds):
    h, w = img.shape[:2]
    px = torch.from_numpy(img.reshape(h // 4, 4, w // 4, 4, 4).transpose(0, 2, 1, 3, 4).reshape(-1, 64).copy()).cuda()
    n = px.shape[0]
    out = torch.zeros((n, 16), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for lvl in levels:
        K.encode_blocks_dev(px.data_ptr(), out.data_ptr(), n, lvl, st)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 3 if lvl < 4 else 1
        e0.record()
        for _ in range(reps):
            K.encode_blocks_dev(px.data_ptr(), out.data_ptr(), n, lvl, st)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        line = "%s %dx%d level %d: %.2f ms  %.1f Mpix/s" % (name, w, h, lvl, ms, w * h / ms / 1e3)
        if R.have_ref() and name == "mix" and n <= 262144:
            sub = slice(0, n, 8)
            want = R.ref_encode_blocks(px.cpu().numpy()[sub], lvl, 16)
            got = out.cpu().numpy()[sub]
            line += "  parity %d/%d" % ((want == got).all(1).sum(), len(want))
        print(line, flush=True)
