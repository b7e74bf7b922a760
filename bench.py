#!/usr/bin/env python3
"""bench.py -- the driver's measurement contract for the UASTC hot path (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|c5]

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path over one batch of synthetic input.

  c2 (default)  BASELINE.json configs[1]: 8192x8192 RGBA8 + full mip chain (14 images, 5 592 407 blocks, 89.5 Mpix)
                encoded to UASTC at level 2, then every block transcoded to BC7_RGBA and to ASTC_4x4_RGBA.
                metric = UASTC encode Mpix/s over the whole step; the line also carries the transcode GB/s.
  c5            BASELINE.json configs[4]: batch transcode of 2^26 UASTC blocks to BC7 (the HBM-streaming roofline run).
Under torchrun every rank runs the same per-GPU workload on its own texture (weak scaling, no collective on the
data path: blocks are independent and the host gathers the shards).
"""
import argparse, json, os, subprocess, sys, threading, time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

LEVEL = 2
C2_NAME = "C2: 8192x8192 RGBA8 + full mip chain (14 images, 5592407 blocks, 89.48 Mpix), UASTC level 2, then transcode to BC7_RGBA and ASTC_4x4_RGBA"
C5_NAME = "C5: batch transcode of 1024 UASTC 1024x1024 textures to BC7 (2^26 blocks, 1 GiB in / 1 GiB out per GPU)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


def wait_for_quiet_gpu(index, timeout_s=20.0):
    """Before the warm-up: wait (bounded) until nvidia-smi reports the GPU idle.  Twice this round a C2 run that started right
    after another process's GPU work on the same box measured the encode kernel at 2.1x and 3.0x its usual duration while every
    later kernel of the same run was normal (profiles/r3e_bench_c2_contended.json): something else still held the GPU.  Waiting
    for an idle GPU skips no work; the wait is reported in config.gpu_quiet_wait_s.  Our own work must be synchronised first."""
    t0 = time.perf_counter()
    time.sleep(0.5)
    quiet = 0
    while time.perf_counter() - t0 < timeout_s:
        try:
            o = subprocess.run(["nvidia-smi", "-i", str(index), "--query-gpu=utilization.gpu", "--format=csv,noheader,nounits"],
                               capture_output=True, text=True, timeout=5).stdout.strip()
            quiet = quiet + 1 if (o.isdigit() and int(o) <= 5) else 0
        except Exception:
            break
        if quiet >= 2:
            break
        time.sleep(0.5)
    return time.perf_counter() - t0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop, self.th = index, [], threading.Event(), None

    def _run(self):
        while not self.stop.is_set():
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([x.strip() for x in o.split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def mip_chain(base):
    """Full chain by 2x2 box average, integer (a+b+c+d+2)>>2; odd sizes never occur for 8192."""
    out = [base]
    while out[-1].shape[0] > 1 or out[-1].shape[1] > 1:
        a = out[-1].astype(np.uint16)
        h, w = a.shape[:2]
        if h > 1 and w > 1:
            m = (a[0::2, 0::2] + a[1::2, 0::2] + a[0::2, 1::2] + a[1::2, 1::2] + 2) >> 2
        elif h > 1:
            m = (a[0::2] + a[1::2] + 1) >> 1
        else:
            m = (a[:, 0::2] + a[:, 1::2] + 1) >> 1
        out.append(m.astype(np.uint8))
    return out


def c2_images(size, seed):
    import refutil as R
    return mip_chain(R.img_mix(size, size, seed))


def n_blocks_of(img):
    return ((img.shape[0] + 3) // 4) * ((img.shape[1] + 3) // 4)


def make_uastc_corpus(n_base_images=4, size=512, level=2):
    """Reference-encoded UASTC blocks of img_mix textures (SURVEY 8(d) C5 recipe, scaled): realistic mode mix."""
    import refutil as R
    if R.have_ref():
        blocks = [R.ref_encode_blocks(R.ref_extract_blocks(R.img_mix(size, size, 200 + j)), level, os.cpu_count() or 8)
                  for j in range(n_base_images)]
        return np.concatenate(blocks), "img_mix %dx%d x%d encoded by the reference at level %d" % (size, size, n_base_images, level)
    g = np.load(os.path.join(ROOT, "tests", "golden", "transcode_golden.npz"))
    return g["uastc"], "tests/golden/transcode_golden.npz corpus (oracle/_ref absent)"


def cpu_reference_sample(size, seed, steps, warmup):
    """The unmodified reference (oracle/_ref) on the host cores: encode a size x size crop of the C2 base level at
    level 2 with all threads, then transcode those blocks to BC7 and ASTC.  Returns dict or None."""
    import refutil as R
    if not R.have_ref():
        return None
    cores = os.cpu_count() or 1
    img = R.img_mix(size, size, seed)
    px = R.ref_extract_blocks(img)
    n = px.shape[0]
    f = R.ref().ref_transcode_mt
    enc_s, xc_s = [], []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        u = R.ref_encode_blocks(px, LEVEL, cores)
        t1 = time.perf_counter()
        o = np.empty_like(u)
        f(u.ctypes.data_as(R._u8p), n, 0, o.ctypes.data_as(R._u8p), cores)
        f(u.ctypes.data_as(R._u8p), n, 1, o.ctypes.data_as(R._u8p), cores)
        t2 = time.perf_counter()
        if it >= warmup:
            enc_s.append(t1 - t0); xc_s.append(t2 - t1)
    e, x = sum(enc_s) / len(enc_s), sum(xc_s) / len(xc_s)
    return {"mpix_s": size * size / (e + x) / 1e6, "encode_mpix_s": size * size / e / 1e6, "transcode_gbs": 2 * 32.0 * n / x / 1e9,
            "step_s": e + x, "cores": cores,
            "sample": "%dx%d img_mix crop (%d blocks): reference encode_uastc level 2 over %d threads + reference transcode to BC7 and ASTC" % (size, size, n, cores)}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref, unmodified sources), all host
    threads, on a bounded sample of the same workload per step."""
    import refutil as R
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if not R.have_ref():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_uastc.so not present"}))
        return 0
    if args.workload == "c5":
        return run_reference_c5(args)
    r = cpu_reference_sample(args.cpu_sample, 2, max(1, args.steps), max(0, min(args.warmup, 1)))
    line = {"impl": "reference", "metric": "uastc_encode_Mpix_per_s", "value": r["mpix_s"], "unit": "Mpix/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["step_s"] * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": C2_NAME, "sample": r["sample"], "level": LEVEL},
            "encode_only_mpix_s": r["encode_mpix_s"], "transcode": {"value": r["transcode_gbs"], "unit": "GB/s"},
            "cpu_baseline": {"value": r["mpix_s"], "unit": "Mpix/s", "cores": r["cores"], "kind": "reference", "sample": r["sample"]},
            "e2e": {"value": r["mpix_s"], "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def run_reference_c5(args):
    import refutil as R
    cores = os.cpu_count() or 1
    base, desc = make_uastc_corpus()
    n = 1 << 22  # bounded sample: 4 Mi blocks = 64 MiB in, 64 MiB out per step
    rng = np.random.default_rng(1)
    blocks = np.ascontiguousarray(base[rng.integers(0, len(base), n)])
    out = np.empty_like(blocks)
    f = R.ref().ref_transcode_mt
    p_in, p_out = blocks.ctypes.data_as(R._u8p), out.ctypes.data_as(R._u8p)
    for _ in range(args.warmup):
        f(p_in, n, 0, p_out, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        f(p_in, n, 0, p_out, cores)
    dt = (time.perf_counter() - t0) / args.steps
    gbs = 32.0 * n / dt / 1e9
    line = {"impl": "reference", "metric": "uastc_to_bc7_transcode_GBps", "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": C5_NAME, "sample": "%d blocks per step" % n, "corpus": desc},
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "reference",
                             "sample": "%d UASTC blocks -> BC7 per step, ref_transcode_mt over %d threads" % (n, cores)},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def dist_env():
    return int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))


def run_c5(args):
    """HBM-streaming roofline run: 2^26 device-resident UASTC blocks -> BC7."""
    import torch
    import torch.distributed as dist
    import ktx_software_b200 as K
    import refutil as R

    world, rank, local = dist_env()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    K.init()
    dev = torch.device("cuda", local)
    pk, pk_kind = peaks()
    n = args.blocks
    base, desc = make_uastc_corpus()
    g = torch.Generator(device="cpu").manual_seed(1000 + rank)
    d_base = torch.from_numpy(base).to(dev)
    idx = torch.randint(0, d_base.shape[0], (n,), generator=g).to(dev)
    d_src = d_base[idx].contiguous()          # [n,16] u8, resident in HBM; 1 GiB >> 126 MB L2
    del idx
    d_dst = torch.empty_like(d_src)
    d_bad = torch.zeros(1, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream()

    def step():
        K.transcode_dev(K.TTF_BC7_RGBA, d_src.data_ptr(), d_dst.data_ptr(), n, stream.cuda_stream, d_bad.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    torch.cuda.synchronize()
    quiet_wait = wait_for_quiet_gpu(local)
    W = max(3, args.warmup)
    for _ in range(W):
        step()
    barrier()
    l0 = K.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local) as cs:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for a, b in evs:
            a.record(); step(); b.record()
        e1.record()
        barrier()
        if cs.rows == []:
            time.sleep(0.3)
    launches = K.launch_count() - l0
    total_ms = e0.elapsed_time(e1)
    kern_ms = sum(a.elapsed_time(b) for a, b in evs) / args.steps
    assert d_bad.item() == -1, "transcode flagged an invalid block"
    t = torch.tensor([total_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = t.item() / args.steps
    value = 32.0 * n * world / (ms_per_step * 1e-3) / 1e9
    parity = None
    if R.have_ref():
        sub = slice(0, 1 << 16)
        parity = bool((R.ref_transcode(d_src[sub].cpu().numpy(), 0)[0] == d_dst[sub].cpu().numpy()).all())

    # e2e: same metric through the C-ABI host call, pinned host buffers, H2D + D2H inside the timed region
    ne = args.e2e_blocks
    h_src = torch.empty((ne, 16), dtype=torch.uint8).pin_memory()
    h_src.copy_(d_src[:ne].cpu())
    h_dst = torch.empty((ne, 16), dtype=torch.uint8).pin_memory()
    spans = [(0, 0, ne)]
    hs, hd = h_src.numpy(), h_dst.numpy().reshape(-1)
    for _ in range(2):
        K.transcode(hs, K.TTF_BC7_RGBA, spans=spans, dst=hd)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        K.transcode(hs, K.TTF_BC7_RGBA, spans=spans, dst=hd)
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / args.steps
    t = torch.tensor([e2e_s], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = 32.0 * ne * world / t.item() / 1e9
    e2e_ok = bool((h_dst[:1 << 14].numpy() == d_dst[:1 << 14].cpu().numpy()).all())

    # The other block targets on a 2^24-block slice of the same input, device-resident, with the reference
    # transcoder timed beside each on a bounded sample.  bytes = 16 read + 8 or 16 written per block.
    targets = {}
    if rank == 0 and world == 1:   # (single-GPU runs only: the other ranks of a multi-GPU run would idle behind rank 0 here)
        nt = min(n, 1 << 24)
        names = [("astc", K.TTF_ASTC_4x4_RGBA), ("bc1", K.TTF_BC1_RGB), ("bc3", K.TTF_BC3_RGBA), ("bc4", K.TTF_BC4_R), ("bc5", K.TTF_BC5_RG),
                 ("etc1", K.TTF_ETC1_RGB), ("etc2", K.TTF_ETC2_RGBA), ("eac_r11", K.TTF_ETC2_EAC_R11), ("eac_rg11", K.TTF_ETC2_EAC_RG11)]
        for name, tf in names:
            ob = K.target_block_bytes(tf)
            for _ in range(2):
                K.transcode_dev_ex(tf, 0, True, d_src.data_ptr(), d_dst.data_ptr(), nt, stream.cuda_stream, d_bad.data_ptr())
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 5
            a.record()
            for _ in range(reps):
                K.transcode_dev_ex(tf, 0, True, d_src.data_ptr(), d_dst.data_ptr(), nt, stream.cuda_stream, d_bad.data_ptr())
            b.record(); torch.cuda.synchronize()
            ms = a.elapsed_time(b) / reps
            ent = {"gbs": (16.0 + ob) * nt / ms / 1e6, "ms": ms, "blocks": nt, "bytes_per_block": 16 + ob,
                   "frac_of_hbm_peak": (16.0 + ob) * nt / ms / 1e6 / pk["hbm_gbs"]}
            if R.have_ref():
                ns = 1 << 16
                hb = np.ascontiguousarray(d_src[:ns].cpu().numpy())
                got = d_dst.reshape(-1)[:ns * ob].cpu().numpy().reshape(ns, ob)
                if tf == K.TTF_ASTC_4x4_RGBA:
                    want = R.ref_transcode(hb, 1)[0]
                else:
                    want = R.ref_transcode_hinted(hb, tf)[0]
                ent["parity_vs_reference"] = bool((got == want).all())
                if tf != K.TTF_ASTC_4x4_RGBA:
                    cores = os.cpu_count() or 1
                    nb = 1 << 20
                    big = np.ascontiguousarray(d_src[:nb].cpu().numpy())
                    chunks = np.array_split(big, cores)
                    from concurrent.futures import ThreadPoolExecutor
                    with ThreadPoolExecutor(cores) as ex:
                        list(ex.map(lambda c: R.ref_transcode_hinted(c, tf), chunks))   # warm
                        t0 = time.perf_counter()
                        list(ex.map(lambda c: R.ref_transcode_hinted(c, tf), chunks))
                        dt = time.perf_counter() - t0
                    ent["cpu_reference_gbs"] = (16.0 + ob) * nb / dt / 1e9
                    ent["cpu_reference_sample"] = "%d blocks, transcode_image over %d threads" % (nb, cores)
            targets[name] = ent
        d_bad.fill_(-1)

    if rank == 0:
        cpu = None
        if R.have_ref():
            cores = os.cpu_count() or 1
            nb = 1 << 22
            hb = np.ascontiguousarray(d_src[:nb].cpu().numpy())
            ho = np.empty_like(hb)
            f = R.ref().ref_transcode_mt
            f(hb.ctypes.data_as(R._u8p), nb, 0, ho.ctypes.data_as(R._u8p), cores)
            t0 = time.perf_counter()
            reps = 3
            for _ in range(reps):
                f(hb.ctypes.data_as(R._u8p), nb, 0, ho.ctypes.data_as(R._u8p), cores)
            dt = (time.perf_counter() - t0) / reps
            cpu = {"value": 32.0 * nb / dt / 1e9, "unit": "GB/s", "cores": cores, "kind": "reference",
                   "sample": "%d blocks -> BC7, unmodified reference transcoder over %d threads, %d reps" % (nb, cores, reps)}
        achieved = 32.0 * n / (kern_ms * 1e-3) / 1e9
        line = {"metric": "uastc_to_bc7_transcode_GBps", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
                "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": C5_NAME, "blocks_per_gpu": n, "corpus": desc,
                           "l2": "inputs (1 GiB) larger than L2, no flush needed", "parity_vs_reference": parity,
                           "gpu_quiet_wait_s": round(quiet_wait, 2)},
                "clocks": cs.summary(),
                "e2e": {"value": e2e_val, "unit": "GB/s", "h2d_bytes_per_step": 16 * ne, "d2h_bytes_per_step": 16 * ne,
                        "blocks": ne, "api": "b200_uastc_transcode (host pinned buffers)", "matches_device_path": e2e_ok},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                             "frac": achieved / pk["hbm_gbs"],
                             # ncu --set full of this workload (profiles/r2i_transcode_tiled_ncu_summary.txt), 2^26 blocks
                             "traffic": (1.076825e9 + 1.027500e9) if n == (1 << 26) else None, "peak_kind": pk_kind,
                             "kernel": "k_transcode_tiled<BC7>", "kernel_ms": kern_ms, "algorithmic_bytes_per_launch": 32 * n},
                "other_targets": targets,
                "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_c2(args):
    """Default workload: encode a full 8192^2 mip chain at level 2, then transcode every block to BC7 and ASTC."""
    import torch
    import torch.distributed as dist
    import ktx_software_b200 as K
    import refutil as R

    world, rank, local = dist_env()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    K.init()
    dev = torch.device("cuda", local)
    pk, pk_kind = peaks()

    imgs = c2_images(args.size, 2 + rank)                      # per-rank texture (weak scaling)
    nblk = [n_blocks_of(im) for im in imgs]
    offs = np.concatenate([[0], np.cumsum(nblk)]).astype(np.int64)
    total_blocks = int(offs[-1])
    total_pix = sum(im.shape[0] * im.shape[1] for im in imgs)
    d_imgs = [torch.from_numpy(im).to(dev) for im in imgs]     # 358 MB resident in HBM (> 126 MB L2)
    d_uastc = torch.zeros((total_blocks, 16), dtype=torch.uint8, device=dev)
    d_bc7 = torch.empty_like(d_uastc)
    d_astc = torch.empty_like(d_uastc)
    d_bad = torch.zeros(1, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream()
    st = stream.cuda_stream
    ev = {k: (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for k in ("enc0", "bc7", "astc")}
    acc = {k: 0.0 for k in ev}

    dev_imgs = [(d.data_ptr(), im.shape[1], im.shape[0], d_uastc.data_ptr() + int(offs[i]) * 16) for i, (im, d) in enumerate(zip(imgs, d_imgs))]

    def step(timed=False):
        if timed:
            ev["enc0"][0].record()
        K.encode_images_dev(dev_imgs, LEVEL, st)          # the whole mip chain in one launch
        if timed:
            ev["enc0"][1].record()
        if timed:
            ev["bc7"][0].record()
        K.transcode_dev(K.TTF_BC7_RGBA, d_uastc.data_ptr(), d_bc7.data_ptr(), total_blocks, st, d_bad.data_ptr())
        if timed:
            ev["bc7"][1].record(); ev["astc"][0].record()
        K.transcode_dev(K.TTF_ASTC_4x4_RGBA, d_uastc.data_ptr(), d_astc.data_ptr(), total_blocks, st, 0)
        if timed:
            ev["astc"][1].record()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    torch.cuda.synchronize()
    quiet_wait = wait_for_quiet_gpu(local)
    W = max(3, args.warmup)
    for _ in range(W):
        step()
    barrier()
    l0 = K.launch_count()
    with ClockSampler(local) as cs:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step(timed=True)
            # per-kernel events are read after the step's work is queued; the sync here is outside any kernel
            ev["astc"][1].synchronize()
            for k in ev:
                acc[k] += ev[k][0].elapsed_time(ev[k][1])
        e1.record()
        barrier()
    launches = K.launch_count() - l0
    assert d_bad.item() == -1, "transcode flagged an invalid block"
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = t.item() / args.steps
    value = total_pix * world / (ms_per_step * 1e-3) / 1e6
    enc0_ms, bc7_ms, astc_ms = (acc[k] / args.steps for k in ("enc0", "bc7", "astc"))

    # parity spot check of this very output against the reference (not timed): a strided sample of the base level
    parity = None
    if R.have_ref():
        h, w = imgs[0].shape[:2]
        bx = w // 4
        rows = np.arange(0, h // 4, max(1, (h // 4) // 16))[:16]
        sel = (rows[:, None] * bx + np.arange(bx)[None, :]).reshape(-1)
        px = np.stack([imgs[0][r * 4:r * 4 + 4].reshape(4, bx, 4, 4).transpose(1, 0, 2, 3).reshape(bx, 64) for r in rows]).reshape(-1, 64)
        got = d_uastc[torch.from_numpy(sel).to(dev)].cpu().numpy()
        want = R.ref_encode_blocks(px, LEVEL, os.cpu_count() or 8)
        ok_e = bool((got == want).all())
        ok_b = bool((R.ref_transcode(got, 0)[0] == d_bc7[torch.from_numpy(sel).to(dev)].cpu().numpy()).all())
        ok_a = bool((R.ref_transcode(got, 1)[0] == d_astc[torch.from_numpy(sel).to(dev)].cpu().numpy()).all())
        parity = {"encode_bit_exact": ok_e, "bc7_bit_exact": ok_b, "astc_bit_exact": ok_a, "blocks_checked": int(len(sel))}

    # ---- e2e: the same step through the host-buffer C ABI (b200_uastc_encode + b200_uastc_transcode x2), pinned
    # host memory, every H2D / D2H copy inside the timed region
    h_imgs = []
    for im in imgs:
        tbuf = torch.empty(im.shape, dtype=torch.uint8).pin_memory()
        tbuf.copy_(torch.from_numpy(im))
        h_imgs.append(tbuf.numpy())
    h_u = torch.empty((total_blocks, 16), dtype=torch.uint8).pin_memory()
    h_b = torch.empty((total_blocks, 16), dtype=torch.uint8).pin_memory()
    h_a = torch.empty((total_blocks, 16), dtype=torch.uint8).pin_memory()
    import ctypes
    arr = (K.api.Image * len(imgs))()
    optr = (ctypes.c_void_p * len(imgs))()
    for i, im in enumerate(h_imgs):
        arr[i] = K.api.Image(im.ctypes.data, im.shape[1], im.shape[0], 0)
        optr[i] = h_u.data_ptr() + int(offs[i]) * 16
    params = K.api.make_params(LEVEL)
    any_alpha = ctypes.c_uint32(0)
    spans = [(0, 0, total_blocks)]
    hu, hb, ha = h_u.numpy().reshape(-1), h_b.numpy().reshape(-1), h_a.numpy().reshape(-1)

    e2e_parts = [0.0, 0.0, 0.0]

    def e2e_step():
        t0 = time.perf_counter()
        K.api._check(K.lib().b200_uastc_encode(arr, len(imgs), ctypes.byref(params), optr, ctypes.byref(any_alpha)))
        t1 = time.perf_counter()
        K.transcode(hu, K.TTF_BC7_RGBA, spans=spans, dst=hb)
        t2 = time.perf_counter()
        K.transcode(hu, K.TTF_ASTC_4x4_RGBA, spans=spans, dst=ha)
        t3 = time.perf_counter()
        for i, d in enumerate((t1 - t0, t2 - t1, t3 - t2)):
            e2e_parts[i] += d

    e2e_step()
    barrier()
    e2e_steps = max(1, min(args.steps, 3))
    e2e_parts[:] = [0.0, 0.0, 0.0]
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    t = torch.tensor([e2e_s], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = total_pix * world / t.item() / 1e6
    e2e_ok = bool((h_u.numpy() == d_uastc.cpu().numpy()).all() and (h_b.numpy() == d_bc7.cpu().numpy()).all()
                  and (h_a.numpy() == d_astc.cpu().numpy()).all())
    in_bytes = sum(im.nbytes for im in imgs)

    if rank == 0:
        cpu = None
        r = cpu_reference_sample(args.cpu_sample, 2, 1, 0)
        if r:
            cpu = {"value": r["mpix_s"], "unit": "Mpix/s", "cores": r["cores"], "kind": "reference", "sample": r["sample"],
                   "encode_only_mpix_s": r["encode_mpix_s"], "transcode_gbs": r["transcode_gbs"]}
        enc_bytes = 80.0 * total_blocks
        xc_bytes = 32.0 * total_blocks
        # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture of this very workload
        # (profiles/r2o_c2_full_size_ncu_summary.txt); only valid for the full-size run
        enc_traffic = (11.586723e9 + 32.283820e9) if args.size == 8192 else None
        bc7_traffic = (89.870592e6 + 44.263680e6) if args.size == 8192 else None
        line = {"metric": "uastc_encode_Mpix_per_s", "value": value, "unit": "Mpix/s", "n_gpus": world, "steps": args.steps,
                "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic",
                "config": {"workload": C2_NAME if args.size == 8192 else "C2 scaled to %d^2" % args.size, "level": LEVEL,
                           "blocks_per_gpu": total_blocks, "pixels_per_gpu": total_pix,
                           "l2": "inputs (358 MB RGBA8 per step) larger than the 126 MB L2, no flush needed",
                           "gpu_quiet_wait_s": round(quiet_wait, 2),
                           "parity_vs_reference": parity},
                "clocks": cs.summary(),
                "e2e": {"value": e2e_val, "unit": "Mpix/s", "h2d_bytes_per_step": in_bytes + 2 * 16 * total_blocks,
                        "d2h_bytes_per_step": 3 * 16 * total_blocks, "steps": e2e_steps,
                        "api": "b200_uastc_encode + b200_uastc_transcode(BC7) + b200_uastc_transcode(ASTC), pinned host buffers",
                        "matches_device_path": e2e_ok,
                        "ms_per_call": {"encode": 1e3 * e2e_parts[0] / e2e_steps, "transcode_bc7": 1e3 * e2e_parts[1] / e2e_steps,
                                        "transcode_astc": 1e3 * e2e_parts[2] / e2e_steps}},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": enc_bytes / (enc0_ms * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                             "frac": enc_bytes / (enc0_ms * 1e-3) / 1e9 / pk["hbm_gbs"], "traffic": enc_traffic, "peak_kind": pk_kind,
                             "kernel": "k_encode_images<21> (all 14 mips in one launch: %d blocks)" % total_blocks, "kernel_ms": enc0_ms,
                             "algorithmic_bytes_per_launch": enc_bytes,
                             "note": "dominant kernel of the step; ALU-issue bound by design (no dense contraction): issue slots 52% "
                                     "busy, 26 of 32 lanes active (profiles/r2m_encode_l2_ncu_summary.txt); HBM fraction reported for "
                                     "completeness; the DRAM traffic above the algorithmic bytes is per-thread local memory (candidate "
                                     "lists, spills at the 64-register cap of a 1024-thread CTA) falling out of L2"},
                "transcode": {"bc7_gbs": xc_bytes / (bc7_ms * 1e-3) / 1e9, "astc_gbs": xc_bytes / (astc_ms * 1e-3) / 1e9,
                              "bc7_ms": bc7_ms, "astc_ms": astc_ms, "blocks": total_blocks,
                              "roofline": {"bound": "hbm", "achieved": xc_bytes / (bc7_ms * 1e-3) / 1e9, "peak": pk["hbm_gbs"],
                                           "unit": "GB/s", "frac": xc_bytes / (bc7_ms * 1e-3) / 1e9 / pk["hbm_gbs"], "traffic": bc7_traffic,
                                           "kernel": "k_transcode_tiled<BC7>", "algorithmic_bytes_per_launch": xc_bytes}},
                "encode_only_mpix_s": total_pix / (enc0_ms * 1e-3) / 1e6,
                "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="c2", choices=["c2", "c5"])
    ap.add_argument("--size", type=int, default=8192, help="c2: base level edge (8192 = the named config)")
    ap.add_argument("--cpu-sample", type=int, default=2048, help="edge of the crop the CPU reference encodes per step")
    ap.add_argument("--blocks", type=int, default=1 << 26)
    ap.add_argument("--e2e-blocks", type=int, default=1 << 25)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    return run_c5(args) if args.workload == "c5" else run_c2(args)


if __name__ == "__main__":
    sys.exit(main())
