#!/usr/bin/env python3
"""bench.py -- the driver's measurement contract for the UASTC hot path (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|c1|c3|c4|c5] [--no-configs]

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path over one batch of synthetic input.

  c2 (default)  BASELINE.json configs[1]: img_mix(8192,8192,2) + 13 mips (14 images, 5 592 407 blocks, 89.5 Mpix) encoded to UASTC
                at level 2, then every block transcoded to BC7_RGBA and to ASTC_4x4_RGBA.  metric = UASTC encode Mpix/s over the
                whole step.  The default run also carries one record per other named config under "configs":
                c1 (2048^2 level 2), c3_l3 / c3_l4_rdo (4096^2: level 3 no RDO; level 4 + RDO lambda 1), c4 (64 x 6 x 2048^2 level 2,
                ONE job sharded over the GPUs by the library's planner), c5 (2^26 blocks -> BC7, the HBM-streaming roofline run).
  c1|c3|c4|c5   that config alone as the headline line.

Multi-GPU (torchrun, one rank per GPU): the headline is weak scaling -- every rank runs the workload on its own texture, no
collective on the data path (blocks are independent), max over ranks of the device time.  The sharded-job record ("configs.c4",
"sharded_job") is strong scaling: rank 0 hands ONE job to the library with device_mask = all N GPUs; the library's planner
cuts it into slabs / images, runs one host thread + lane per GPU and gathers into the caller's buffer (other ranks idle at a
CPU barrier meanwhile).
"""
import argparse, ctypes, json, os, subprocess, sys, threading, time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

LEVEL = 2
C1_NAME = "C1: img_mix(2048,2048,1) RGBA8, no mips (262144 blocks, 4.19 Mpix), UASTC level 2"
C2_NAME = "C2: 8192x8192 RGBA8 + full mip chain (14 images, 5592407 blocks, 89.48 Mpix), UASTC level 2, then transcode to BC7_RGBA and ASTC_4x4_RGBA"
C3_NAME = "C3: img_mix(4096,4096,3) RGBA8 (1048576 blocks, 16.8 Mpix): UASTC level 3 without RDO (bit-exact tier) and level 4 + RDO lambda=1.0 (dict 4096, 4 jobs)"
C4_NAME = "C4: cubemap array 64 layers x 6 faces x 2048^2 RGBA8 (384 images, 100663296 blocks, 1610.6 Mpix), UASTC level 2, ONE job sharded by face over the GPUs"
C5_NAME = "C5: batch transcode of 1024 UASTC 1024x1024 textures to BC7 (2^26 blocks, 1 GiB in / 1 GiB out per GPU)"
SM_ISSUE_PER_CLK = 148 * 4   # warp-instructions per clock, whole GPU (one per SM sub-partition)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "sm_max_mhz": 1965.0}, "fallback"


def ncu_counters(build):
    """Per-kernel counters (warp instructions, DRAM bytes) from the committed ncu capture of THIS build (tools/ncu_counters.py
    writes profiles/ncu_counters.json keyed by the library's build id).  No entry for the loaded build -> None: a counter
    measured on another build is never quoted."""
    p = os.path.join(ROOT, "profiles", "ncu_counters.json")
    if not os.path.exists(p):
        return None
    return json.load(open(p)).get(build)


class ClockSampler:
    """SM clock and throttle reasons of one GPU sampled DURING the timed region, in-process through NVML (a subprocess per
    sample per rank saturated the driver lock at 8 ranks).  Falls back to one nvidia-smi query per second."""

    def __init__(self, index):
        self.index, self.rows, self.stop, self.th, self.h = index, [], threading.Event(), None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = index
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                idx = int(vis.split(",")[index])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nv = pynvml
        except Exception:
            self.h = None

    def _sample(self):
        if self.h is not None:
            nv = self.nv
            sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
            mx = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            names = []
            for bit, name in ((nv.nvmlClocksThrottleReasonHwSlowdown, "hw_slowdown"), (nv.nvmlClocksThrottleReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                              (nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_thermal_slowdown"), (nv.nvmlClocksThrottleReasonSwPowerCap, "sw_power_cap")):
                if r & bit:
                    names.append(name)
            return sm, mx, names
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                           capture_output=True, text=True, timeout=5).stdout.strip().split(",")
        names = [n for n, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], o[2:]) if v.strip().lower().startswith("active")]
        return int(o[0]), int(o[1]), names

    def _run(self):
        while not self.stop.is_set():
            try:
                self.rows.append(self._sample())
            except Exception:
                pass
            self.stop.wait(0.02 if self.h is not None else 1.0)

    def __enter__(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        sm = sorted(r[0] for r in self.rows)
        mx = [r[1] for r in self.rows]
        reasons = sorted({n for r in self.rows for n in r[2]})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(self.rows),
                "how": "NVML in-process, 50 Hz" if self.h is not None else "nvidia-smi, 1 Hz"}


def mip_chain(base):
    """Full chain by 2x2 box average, integer (a+b+c+d+2)>>2; odd sizes never occur for powers of two."""
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


def n_blocks_of(img):
    return ((img.shape[0] + 3) // 4) * ((img.shape[1] + 3) // 4)


def blocks_of_rows(img, block_rows):
    """[n,64] texel blocks of the given block rows of an image whose width is a multiple of 4 (raster order inside the row)."""
    bx = img.shape[1] // 4
    return np.stack([img[r * 4:r * 4 + 4].reshape(4, bx, 4, 4).transpose(1, 0, 2, 3).reshape(bx, 64) for r in block_rows]).reshape(-1, 64)


def c2_sample_rows(size):
    """The block rows of the C2 base level both arms' CPU legs encode: every 16th block row, i.e. 1/16 of the level spread
    evenly over the image (content differs across it: alpha only in the left half, solid and grey tiles every few tiles)."""
    return np.arange(0, size // 4, 16)


def make_uastc_corpus(n_base_images=4, size=512, level=2):
    """Reference-encoded UASTC blocks of img_mix textures (SURVEY 8(d) C5 recipe, scaled): realistic mode mix."""
    import refutil as R
    if R.have_ref():
        blocks = [R.ref_encode_blocks(R.ref_extract_blocks(R.img_mix(size, size, 200 + j)), level, os.cpu_count() or 8)
                  for j in range(n_base_images)]
        return np.concatenate(blocks), "img_mix %dx%d x%d encoded by the reference at level %d" % (size, size, n_base_images, level)
    g = np.load(os.path.join(ROOT, "tests", "golden", "transcode_golden.npz"))
    return g["uastc"], "tests/golden/transcode_golden.npz corpus (oracle/_ref absent)"


def ref_encode_transcode(px, flags, steps=1, warmup=0):
    """The unmodified reference (oracle/_ref) on all host cores: encode_uastc over px [n,64] with `flags`, then transcode the result to
    BC7 and ASTC.  -> dict(encode_s, transcode_s, blocks [n,16], cores)."""
    import refutil as R
    cores = os.cpu_count() or 1
    n = px.shape[0]
    f = R.ref().ref_transcode_mt
    enc_s, xc_s, u = [], [], None
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        u = R.ref_encode_blocks(px, flags, cores)
        t1 = time.perf_counter()
        o = np.empty_like(u)
        f(u.ctypes.data_as(R._u8p), n, 0, o.ctypes.data_as(R._u8p), cores)
        f(u.ctypes.data_as(R._u8p), n, 1, o.ctypes.data_as(R._u8p), cores)
        t2 = time.perf_counter()
        if it >= warmup:
            enc_s.append(t1 - t0); xc_s.append(t2 - t1)
    return {"encode_s": sum(enc_s) / len(enc_s), "transcode_s": sum(xc_s) / len(xc_s), "blocks": u, "cores": cores}


def c2_reference_sample(size, seed, steps, warmup, base=None):
    import refutil as R
    if not R.have_ref():
        return None
    if base is None:
        base = R.img_mix(size, size, seed)
    rows = c2_sample_rows(size)
    px = blocks_of_rows(base, rows)
    r = ref_encode_transcode(px, LEVEL, steps, warmup)
    n = px.shape[0]
    e, x = r["encode_s"], r["transcode_s"]
    return {"mpix_s": 16 * n / (e + x) / 1e6, "encode_mpix_s": 16 * n / e / 1e6, "transcode_gbs": 2 * 32.0 * n / x / 1e9, "step_s": e + x,
            "cores": r["cores"], "blocks": r["blocks"], "rows": rows,
            "sample": "block rows r = 0 mod 16 of the %dx%d base level (%d block rows x %d blocks = %d blocks, %.2f Mpix per step): reference encode_uastc "
                      "level 2 over %d threads (nproc = %d) + reference transcode of those blocks to BC7 and ASTC" % (size, size, len(rows), size // 4, n, 16 * n / 1e6, r["cores"], r["cores"])}


# ---------------------------------------------------------------------------------------------------------
def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref, unmodified sources), all host
    threads, on a bounded sample of the SAME image the GPU arm encodes, per step."""
    import refutil as R
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if not R.have_ref():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_uastc.so not present"}))
        return 0
    cores = os.cpu_count() or 1
    wl = args.workload
    common = {"impl": "reference", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
              "vs_baseline": None, "dtype": "u8", "data": "synthetic"}
    if wl == "c5":
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
        v, unit, metric, ms = 32.0 * n / dt / 1e9, "GB/s", "uastc_to_bc7_transcode_GBps", dt * 1e3
        cfg = {"workload": C5_NAME, "sample": "%d UASTC blocks -> BC7 per step, ref_transcode_mt over %d threads (nproc = %d)" % (n, cores, cores), "corpus": desc}
    elif wl in ("c1", "c3", "c4"):
        if wl == "c1":
            img, flags, name, samp = R.img_mix(2048, 2048, 1), 2, C1_NAME, "the whole image (262144 blocks)"
            px = R.ref_extract_blocks(img)
        elif wl == "c4":
            img, flags, name, samp = R.img_mix(2048, 2048, 100), 2, C4_NAME, "face 0 of 384 (262144 blocks); faces are independent, so Mpix/s carries over"
            px = R.ref_extract_blocks(img)
        else:
            img, flags, name = R.img_mix(4096, 4096, 3), 3, C3_NAME
            rows = np.arange(0, 1024, 8)
            px = blocks_of_rows(img, rows)
            samp = "level 3 without RDO on block rows r = 0 mod 8 (131072 blocks)"
        ts = []
        for it in range(max(0, min(args.warmup, 1)) + max(1, args.steps)):
            t0 = time.perf_counter()
            R.ref_encode_blocks(px, flags, cores)
            if it >= max(0, min(args.warmup, 1)):
                ts.append(time.perf_counter() - t0)
        dt = sum(ts) / len(ts)
        v, unit, metric, ms = 16 * px.shape[0] / dt / 1e6, "Mpix/s", "uastc_encode_Mpix_per_s", dt * 1e3
        cfg = {"workload": name, "sample": samp + ": reference encode_uastc over %d threads (nproc = %d)" % (cores, cores)}
    else:
        r = c2_reference_sample(args.size, 2, max(1, args.steps), max(0, min(args.warmup, 1)))
        v, unit, metric, ms = r["mpix_s"], "Mpix/s", "uastc_encode_Mpix_per_s", r["step_s"] * 1e3
        cfg = {"workload": C2_NAME if args.size == 8192 else "C2 scaled to %d^2" % args.size, "sample": r["sample"], "level": LEVEL}
        common.update({"encode_only_mpix_s": r["encode_mpix_s"], "transcode": {"value": r["transcode_gbs"], "unit": "GB/s"}})
    line = dict(common)
    line.update({"metric": metric, "value": v, "unit": unit, "ms_per_step": ms, "config": cfg,
                 "cpu_baseline": {"value": v, "unit": unit, "cores": cores, "kind": "reference", "sample": cfg["sample"]},
                 "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------------
class Ctx:
    def __init__(self, args):
        import torch
        import torch.distributed as dist
        import ktx_software_b200 as K
        import refutil as R
        self.args, self.torch, self.dist, self.K, self.R = args, torch, dist, K, R
        self.world, self.rank, self.local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.cpu_group = None
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.cpu_group = dist.new_group(backend="gloo")   # barriers that must not put a spinning kernel on the GPUs
        K.init()
        self.dev = torch.device("cuda", self.local)
        self.pk, self.pk_kind = peaks()
        self.build = K.build_id()
        self.counters = ncu_counters(self.build)
        self.stream = torch.cuda.current_stream()
        self.st = self.stream.cuda_stream
        self.sm_max_mhz = float(self.pk.get("sm_max_mhz", 1965.0))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def cpu_barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier(group=self.cpu_group)

    def max_over_ranks(self, v):
        t = self.torch.tensor([float(v)], device=self.dev, dtype=self.torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.item()

    def gather(self, v):
        t = self.torch.tensor([float(v)], device=self.dev, dtype=self.torch.float64)
        if self.world == 1:
            return [t.item()]
        out = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [o.item() for o in out]

    def ev(self):
        return self.torch.cuda.Event(enable_timing=True)

    def pinned(self, arr):
        t = self.torch.empty(arr.shape, dtype=self.torch.uint8).pin_memory()
        t.copy_(self.torch.from_numpy(arr))
        return t

    def issue_roofline(self, kernel_key, kernel_name, blocks, kernel_ms, alg_bytes):
        """Encode kernels are bound by SM instruction issue (no dense contraction, 80 B of HBM traffic per block): achieved =
        warp-instructions of the launch (ncu count of THIS build, per block, x blocks) / kernel time; peak = 148 SMs x 4
        schedulers x max SM clock.  The HBM figure rides along as a secondary field."""
        peak = SM_ISSUE_PER_CLK * self.sm_max_mhz * 1e6 / 1e12
        c = (self.counters or {}).get(kernel_key)
        hbm = alg_bytes / (kernel_ms * 1e-3) / 1e9
        r = {"bound": "sm_issue", "achieved": None, "peak": peak, "unit": "T warp-inst/s", "frac": None, "traffic": None, "kernel": kernel_name, "kernel_ms": kernel_ms,
             "peak_how": "148 SMs x 4 schedulers x %.0f MHz (max SM clock)" % self.sm_max_mhz, "build": self.build,
             "hbm": {"algorithmic_bytes_per_launch": alg_bytes, "achieved_gbs": hbm, "frac_of_hbm_peak": hbm / self.pk["hbm_gbs"], "peak_gbs": self.pk["hbm_gbs"], "peak_kind": self.pk_kind}}
        if c:
            wi = c["warp_inst_per_block"] * blocks
            r["achieved"] = wi / (kernel_ms * 1e-3) / 1e12
            r["frac"] = r["achieved"] / peak
            r["warp_inst_per_block"] = c["warp_inst_per_block"]
            r["counters_from"] = c.get("source")
            if c.get("blocks") == blocks:
                r["traffic"] = c.get("dram_bytes")
            r["issue_active_pct_ncu"] = c.get("issue_active_pct")
        else:
            r["note"] = "no ncu capture of build %s in profiles/ncu_counters.json: instruction count and DRAM traffic not quoted" % self.build
        return r

    def hbm_roofline(self, kernel_key, kernel_name, blocks, kernel_ms, alg_bytes):
        a = alg_bytes / (kernel_ms * 1e-3) / 1e9
        c = (self.counters or {}).get(kernel_key)
        r = {"bound": "hbm", "achieved": a, "peak": self.pk["hbm_gbs"], "unit": "GB/s", "frac": a / self.pk["hbm_gbs"], "traffic": None, "peak_kind": self.pk_kind,
             "kernel": kernel_name, "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": alg_bytes, "build": self.build}
        if c:
            if c.get("blocks") == blocks:
                r["traffic"] = c.get("dram_bytes")
            r["warp_inst_per_block"] = c.get("warp_inst_per_block")
            r["counters_from"] = c.get("source")
        return r


def timed_loop(cx, step, steps, warmup, names):
    """W warm-up steps, then K timed steps between a barrier + synchronize on both sides.  step(evs) records one
    (start, end) event pair per name; nothing is synchronised inside the loop -- the pairs are read afterwards.
    -> (ms per step, max over ranks), {name: ms per launch on this rank}, launches, clocks"""
    K = cx.K
    for _ in range(warmup):
        step(None)
    cx.barrier()
    l0 = K.launch_count()
    evs = [{n: (cx.ev(), cx.ev()) for n in names} for _ in range(steps)]
    with ClockSampler(cx.local) as cs:
        e0, e1 = cx.ev(), cx.ev()
        e0.record()
        for k in range(steps):
            step(evs[k])
        e1.record()
        cx.barrier()
        if not cs.rows:
            time.sleep(0.1)
    launches = K.launch_count() - l0
    per = {n: sum(e[n][0].elapsed_time(e[n][1]) for e in evs) / steps for n in names}
    return cx.max_over_ranks(e0.elapsed_time(e1)) / steps, per, launches, cs.summary()


def e2e_loop(cx, fn, steps):
    """fn() = the step through the host-buffer C ABI (pinned host memory, every H2D / D2H inside).  Wall clock, max over ranks."""
    fn()
    cx.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    cx.torch.cuda.synchronize()
    return cx.max_over_ranks((time.perf_counter() - t0) / steps)


# ---------------------------------------------------------------------------------------------------------
def bench_c2(cx, args):
    """Encode a full 8192^2 mip chain at level 2, then transcode every block to BC7 and ASTC."""
    torch, K, R = cx.torch, cx.K, cx.R
    base = R.img_mix(args.size, args.size, 2 + cx.rank)            # per-rank texture (weak scaling)
    imgs = mip_chain(base)
    nblk = [n_blocks_of(im) for im in imgs]
    offs = np.concatenate([[0], np.cumsum(nblk)]).astype(np.int64)
    total_blocks = int(offs[-1])
    total_pix = sum(im.shape[0] * im.shape[1] for im in imgs)
    d_imgs = [torch.from_numpy(im).to(cx.dev) for im in imgs]     # 358 MB resident in HBM (> 126 MB L2)
    d_uastc = torch.zeros((total_blocks, 16), dtype=torch.uint8, device=cx.dev)
    d_bc7 = torch.empty_like(d_uastc)
    d_astc = torch.empty_like(d_uastc)
    d_bad = torch.zeros(1, dtype=torch.int64, device=cx.dev)
    st = cx.st
    dev_imgs = [(d.data_ptr(), im.shape[1], im.shape[0], d_uastc.data_ptr() + int(offs[i]) * 16) for i, (im, d) in enumerate(zip(imgs, d_imgs))]

    def step(ev):
        if ev: ev["enc"][0].record()
        K.encode_images_dev(dev_imgs, LEVEL, st)          # the whole mip chain in one launch (descriptors by value)
        if ev: ev["enc"][1].record(); ev["bc7"][0].record()
        K.transcode_dev(K.TTF_BC7_RGBA, d_uastc.data_ptr(), d_bc7.data_ptr(), total_blocks, st, d_bad.data_ptr())
        if ev: ev["bc7"][1].record(); ev["astc"][0].record()
        K.transcode_dev(K.TTF_ASTC_4x4_RGBA, d_uastc.data_ptr(), d_astc.data_ptr(), total_blocks, st, 0)
        if ev: ev["astc"][1].record()

    ms_per_step, per, launches, clocks = timed_loop(cx, step, args.steps, max(3, args.warmup), ("enc", "bc7", "astc"))
    assert d_bad.item() == -1, "transcode flagged an invalid block"
    value = total_pix * cx.world / (ms_per_step * 1e-3) / 1e6
    rank_enc_ms = cx.gather(per["enc"])

    # ---- e2e: the same step through the host-buffer C ABI (b200_uastc_encode + b200_uastc_transcode x2), pinned
    # host memory, every H2D / D2H copy inside the timed region
    h_imgs = [cx.pinned(im).numpy() for im in imgs]
    h_u = torch.empty((total_blocks, 16), dtype=torch.uint8).pin_memory()
    h_b = torch.empty((total_blocks, 16), dtype=torch.uint8).pin_memory()
    h_a = torch.empty((total_blocks, 16), dtype=torch.uint8).pin_memory()
    arr = (K.api.Image * len(imgs))()
    optr = (ctypes.c_void_p * len(imgs))()
    for i, im in enumerate(h_imgs):
        arr[i] = K.api.Image(im.ctypes.data, im.shape[1], im.shape[0], 0)
        optr[i] = h_u.data_ptr() + int(offs[i]) * 16
    params = K.api.make_params(LEVEL)
    any_alpha = ctypes.c_uint32(0)
    spans = [(0, 0, total_blocks)]
    hu, hb, ha = h_u.numpy().reshape(-1), h_b.numpy().reshape(-1), h_a.numpy().reshape(-1)
    parts = [0.0, 0.0, 0.0]

    def e2e_step():
        t0 = time.perf_counter()
        K.api._check(K.lib().b200_uastc_encode(arr, len(imgs), ctypes.byref(params), optr, ctypes.byref(any_alpha)))
        t1 = time.perf_counter()
        K.transcode(hu, K.TTF_BC7_RGBA, spans=spans, dst=hb)
        t2 = time.perf_counter()
        K.transcode(hu, K.TTF_ASTC_4x4_RGBA, spans=spans, dst=ha)
        t3 = time.perf_counter()
        for i, d in enumerate((t1 - t0, t2 - t1, t3 - t2)):
            parts[i] += d

    e2e_step()
    parts[:] = [0.0, 0.0, 0.0]
    e2e_s = e2e_loop(cx, e2e_step, args.steps)
    parts = [p / (args.steps + 1) for p in parts]
    e2e_val = total_pix * cx.world / e2e_s / 1e6
    e2e_ok = bool((h_u.numpy() == d_uastc.cpu().numpy()).all() and (h_b.numpy() == d_bc7.cpu().numpy()).all()
                  and (h_a.numpy() == d_astc.cpu().numpy()).all())
    in_bytes = sum(im.nbytes for im in imgs)

    # ---- e2e with the device-resident mip chain: only the base level is uploaded (b200_uastc_encode_mipchain)
    mip = None
    if args.size >= 64:
        base_img = K.api.Image(h_imgs[0].ctypes.data, imgs[0].shape[1], imgs[0].shape[0], 0)
        h_u2 = torch.empty((total_blocks, 16), dtype=torch.uint8).pin_memory()
        optr2 = (ctypes.c_void_p * len(imgs))(*[h_u2.data_ptr() + int(offs[i]) * 16 for i in range(len(imgs))])
        hu2 = h_u2.numpy().reshape(-1)

        def mip_step():
            K.api._check(K.lib().b200_uastc_encode_mipchain(ctypes.byref(base_img), len(imgs), ctypes.byref(params), optr2, None, ctypes.byref(any_alpha)))
            K.transcode(hu2, K.TTF_BC7_RGBA, spans=spans, dst=hb)
            K.transcode(hu2, K.TTF_ASTC_4x4_RGBA, spans=spans, dst=ha)

        mip_s = e2e_loop(cx, mip_step, args.steps)
        mip = {"value": total_pix * cx.world / mip_s / 1e6, "unit": "Mpix/s", "h2d_bytes_per_step": imgs[0].nbytes + 2 * 16 * total_blocks,
               "d2h_bytes_per_step": 3 * 16 * total_blocks, "api": "b200_uastc_encode_mipchain (base level uploaded, 13 levels box-filtered on the device) + 2 x b200_uastc_transcode",
               "matches_host_chain": bool((h_u2.numpy() == h_u.numpy()).all())}

    if cx.rank != 0:
        return None
    # parity of this very output against the reference (not timed), and the CPU baseline, on the same sample of the same image
    parity, cpu = None, None
    r = c2_reference_sample(args.size, 2, 1, 0, base=base)
    if r:
        bx = args.size // 4
        sel = (r["rows"][:, None] * bx + np.arange(bx)[None, :]).reshape(-1)
        tsel = torch.from_numpy(sel).to(cx.dev)
        got = d_uastc[tsel].cpu().numpy()
        ok_e = bool((got == r["blocks"]).all())
        ok_b = bool((R.ref_transcode(got, 0)[0] == d_bc7[tsel].cpu().numpy()).all())
        ok_a = bool((R.ref_transcode(got, 1)[0] == d_astc[tsel].cpu().numpy()).all())
        parity = {"encode_bit_exact": ok_e, "bc7_bit_exact": ok_b, "astc_bit_exact": ok_a, "blocks_checked": int(len(sel)),
                  "full_size_sweep": "profiles/r7_parity_sweep.json (all 5592407 blocks of this config, 0 mismatches)"}
        cpu = {"value": r["mpix_s"], "unit": "Mpix/s", "cores": r["cores"], "kind": "reference", "sample": r["sample"],
               "encode_only_mpix_s": r["encode_mpix_s"], "transcode_gbs": r["transcode_gbs"]}
    enc_key = "k_encode_jobs@c2"
    line = {"metric": "uastc_encode_Mpix_per_s", "value": value, "unit": "Mpix/s", "n_gpus": cx.world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": C2_NAME if args.size == 8192 else "C2 scaled to %d^2" % args.size, "level": LEVEL,
                       "blocks_per_gpu": total_blocks, "pixels_per_gpu": total_pix,
                       "l2": "inputs (358 MB RGBA8 per step) larger than the 126 MB L2, no flush needed",
                       "parity_vs_reference": parity, "build": cx.build},
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": "Mpix/s", "h2d_bytes_per_step": in_bytes + 2 * 16 * total_blocks,
                    "d2h_bytes_per_step": 3 * 16 * total_blocks, "steps": args.steps,
                    "api": "b200_uastc_encode + b200_uastc_transcode(BC7) + b200_uastc_transcode(ASTC), pinned host buffers",
                    "matches_device_path": e2e_ok,
                    "ms_per_call": {"encode": 1e3 * parts[0], "transcode_bc7": 1e3 * parts[1], "transcode_astc": 1e3 * parts[2]},
                    "base_level_only_upload": mip},
            "gpu_launches": launches,
            "roofline": cx.issue_roofline(enc_key, "k_encode_jobs (all 14 mips in one launch: %d blocks)" % total_blocks, total_blocks, per["enc"], 80.0 * total_blocks),
            "kernel_ms_per_rank": {"encode": rank_enc_ms},
            "transcode": {"bc7_gbs": 32.0 * total_blocks / (per["bc7"] * 1e-3) / 1e9, "astc_gbs": 32.0 * total_blocks / (per["astc"] * 1e-3) / 1e9,
                          "bc7_ms": per["bc7"], "astc_ms": per["astc"], "blocks": total_blocks,
                          "roofline": cx.hbm_roofline("k_transcode_tiled<0>@c2", "k_transcode_tiled<BC7>", total_blocks, per["bc7"], 32.0 * total_blocks)},
            "encode_only_mpix_s": total_pix / (per["enc"] * 1e-3) / 1e6,
            "cpu_baseline": cpu}
    return line


def bench_single_image(cx, args, name, img, flags, key, ref_px=None, ref_desc=None, steps=None):
    """One image, no RDO: device-timed k_encode_image, e2e through b200_uastc_encode, parity + CPU baseline on ref_px (default: all blocks)."""
    torch, K, R = cx.torch, cx.K, cx.R
    steps = steps or args.steps
    h, w = img.shape[:2]
    n = n_blocks_of(img)
    d_img = torch.from_numpy(img).to(cx.dev)
    d_out = torch.zeros((n, 16), dtype=torch.uint8, device=cx.dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=cx.dev) if img.nbytes < (160 << 20) else None

    def step(ev):
        if flush is not None:
            flush.fill_(1)          # inputs smaller than the 126 MB L2: flush it between timed iterations
        if ev: ev["enc"][0].record()
        K.encode_image_dev(d_img.data_ptr(), w, h, d_out.data_ptr(), flags, cx.st)
        if ev: ev["enc"][1].record()

    _, per, launches, clocks = timed_loop(cx, step, steps, 3, ("enc",))
    kernel_ms = per["enc"]
    h_img = cx.pinned(img).numpy()
    h_out = torch.empty((n, 16), dtype=torch.uint8).pin_memory()
    arr = (K.api.Image * 1)(K.api.Image(h_img.ctypes.data, w, h, 0))
    optr = (ctypes.c_void_p * 1)(h_out.data_ptr())
    params = K.api.make_params(flags)
    aa = ctypes.c_uint32(0)
    e2e_s = e2e_loop(cx, lambda: K.api._check(K.lib().b200_uastc_encode(arr, 1, ctypes.byref(params), optr, ctypes.byref(aa))), steps)
    rec = {"workload": name, "metric": "uastc_encode_Mpix_per_s", "unit": "Mpix/s", "value": w * h / (kernel_ms * 1e-3) / 1e6, "kernel_ms": kernel_ms, "blocks": n,
           "steps": steps, "l2": "256 MB flush write between timed iterations" if flush is not None else "input larger than L2",
           "e2e": {"value": w * h / e2e_s / 1e6, "unit": "Mpix/s", "h2d_bytes_per_step": img.nbytes, "d2h_bytes_per_step": 16 * n, "api": "b200_uastc_encode, pinned host buffers",
                   "matches_device_path": bool((h_out.numpy() == d_out.cpu().numpy()).all())},
           "gpu_launches": launches, "clocks": clocks,
           "roofline": cx.issue_roofline(key, "k_encode_image<170>" if (flags & 7) >= 4 else "k_encode_jobs", n, kernel_ms, 80.0 * n)}
    if R.have_ref():
        cores = os.cpu_count() or 1
        sel = None
        if ref_px is None:
            ref_px, ref_desc = R.ref_extract_blocks(img), "all %d blocks" % n
        else:
            ref_px, sel = ref_px
        t0 = time.perf_counter()
        want = R.ref_encode_blocks(ref_px, flags, cores)
        dt = time.perf_counter() - t0
        got = d_out.cpu().numpy() if sel is None else d_out.cpu().numpy()[sel]
        rec["parity_vs_reference"] = {"encode_bit_exact": bool((got == want).all()), "blocks_checked": int(len(want))}
        rec["cpu_baseline"] = {"value": 16 * len(want) / dt / 1e6, "unit": "Mpix/s", "cores": cores, "kind": "reference",
                               "sample": "%s of the same image: reference encode_uastc flags %d over %d threads (nproc = %d)" % (ref_desc, flags, cores, cores)}
    return rec


def bench_c1(cx, args):
    return bench_single_image(cx, args, C1_NAME, cx.R.img_mix(2048, 2048, 1), 2, "k_encode_jobs@c1")


def bench_c3(cx, args):
    """C3: level 3 without RDO (bit-exact tier) and level 4 + RDO lambda 1 (PSNR tier)."""
    torch, K, R = cx.torch, cx.K, cx.R
    img = R.img_mix(4096, 4096, 3)
    rows = np.arange(0, 1024, 8)
    sel = (rows[:, None] * 1024 + np.arange(1024)[None, :]).reshape(-1)
    l3 = bench_single_image(cx, args, C3_NAME + " -- level 3, no RDO", img, 3, "k_encode_jobs@c3_l3",
                            ref_px=(blocks_of_rows(img, rows), sel), ref_desc="block rows r = 0 mod 8 (131072 blocks)")
    l3["parity_vs_reference"]["full_size_sweep"] = "profiles/r7_parity_sweep.json (all 1048576 blocks, 0 mismatches)"
    # level 4 + RDO through the host API (the RDO post-pass has no device-pointer image entry): flags 4 | 512 as libktx sets them
    flags = 4 | 512
    n = n_blocks_of(img)
    h_img = cx.pinned(img).numpy()
    h_out = torch.empty((n, 16), dtype=torch.uint8).pin_memory()
    arr = (K.api.Image * 1)(K.api.Image(h_img.ctypes.data, 4096, 4096, 0))
    optr = (ctypes.c_void_p * 1)(h_out.data_ptr())
    params = K.api.make_params(flags, rdo=True, rdo_lambda=1.0, total_jobs=4)
    aa = ctypes.c_uint32(0)
    l0 = K.launch_count()
    t0 = time.perf_counter()
    K.api._check(K.lib().b200_uastc_encode(arr, 1, ctypes.byref(params), optr, ctypes.byref(aa)))
    dt = time.perf_counter() - t0
    # the encode half alone, device-timed
    d_img = torch.from_numpy(img).to(cx.dev)
    d_out = torch.zeros((n, 16), dtype=torch.uint8, device=cx.dev)
    a, b = cx.ev(), cx.ev()
    K.encode_image_dev(d_img.data_ptr(), 4096, 4096, d_out.data_ptr(), flags, cx.st)
    a.record(); K.encode_image_dev(d_img.data_ptr(), 4096, 4096, d_out.data_ptr(), flags, cx.st); b.record()
    torch.cuda.synchronize()
    enc_ms = a.elapsed_time(b)
    rdo = {"workload": C3_NAME + " -- level 4 + RDO", "metric": "uastc_encode_Mpix_per_s", "unit": "Mpix/s", "value": 16 * n / dt / 1e6, "ms_per_step": dt * 1e3, "steps": 1,
           "note": "value == e2e: b200_uastc_encode(rdo_enable) with pinned host buffers; parity mode (the reference's job tiling: 4 ranges + remainder per image, so <= 5 serial chains walk concurrently)",
           "e2e": {"value": 16 * n / dt / 1e6, "unit": "Mpix/s", "h2d_bytes_per_step": img.nbytes, "d2h_bytes_per_step": 16 * n, "api": "b200_uastc_encode(rdo_enable=1, lambda=1, jobs=4)"},
           "encode_level4_kernel_ms": enc_ms, "encode_level4_mpix_s": 16 * n / (enc_ms * 1e-3) / 1e6, "rdo_ms": dt * 1e3 - enc_ms,
           "rdo_us_per_chain_step": (dt * 1e3 - enc_ms) * 1e3 / (n / 4), "gpu_launches": K.launch_count() - l0,
           "roofline": cx.issue_roofline("k_encode_image<170>@c3_l4", "k_encode_image<170> (level 4 half of the step)", n, enc_ms, 80.0 * n)}
    if R.have_ref():
        # parity + CPU baseline on a 1024 x 1024 crop as its own image (RDO job ranges depend on the image, so the crop is run through
        # both implementations whole): reference level-4 encode ~10 s + RDO on 4 job threads ~2 s on 16 cores
        crop = np.ascontiguousarray(img[1536:2560, 1536:2560])
        cores = os.cpu_count() or 1
        ours, _ = K.encode_images([crop], flags, rdo=True, rdo_lambda=1.0, total_jobs=4)
        px = R.ref_extract_blocks(crop)
        t0 = time.perf_counter()
        pre = R.ref_encode_blocks(px, flags, cores)
        t1 = time.perf_counter()
        ref = R.ref_rdo(pre, px, flags, 4, 4, lam=1.0, dict_size=4096)
        t2 = time.perf_counter()

        def psnr(blocks):
            mse = ((R.ref_unpack_rgba(blocks)[0].astype(np.float64) - px.astype(np.float64)) ** 2).mean()
            return 99.0 if mse == 0 else float(10 * np.log10(255.0 ** 2 / mse))

        pa, pb = psnr(ours[0]), psnr(ref)
        rdo["parity_vs_reference"] = {"sample": "rows/cols 1536..2559 of the image as a 1024x1024 image of its own (65536 blocks)", "bit_exact": bool((ours[0] == ref).all()),
                                      "psnr_ours_db": pa, "psnr_reference_db": pb, "psnr_delta_db": abs(pa - pb), "within_0.01_db": bool(abs(pa - pb) <= 0.01),
                                      "mode_agreement": float((R.block_modes(ours[0]) == R.block_modes(ref)).mean()),
                                      "full_size_sweep": "profiles/r7_parity_sweep.json (all 1048576 blocks of this config bit-exact; PSNR, mode agreement, zstd-19 sizes there)"}
        rdo["cpu_baseline"] = {"value": 16 * len(px) / (t2 - t0) / 1e6, "unit": "Mpix/s", "cores": cores, "kind": "reference",
                               "sample": "the same 1024x1024 crop: reference encode_uastc level 4 over %d threads (%.1f s) + uastc_rdo on 4 job threads (%.1f s)" % (cores, t1 - t0, t2 - t1)}
    return l3, rdo


def bench_c4(cx, args, n_images=384):
    """ONE job -- 384 faces of 2048^2 at level 2 -- handed to b200_uastc_encode with device_mask = all GPUs of the run: the
    library's planner spreads faces / slabs over the devices (one host thread + lane each) and gathers into the caller's buffer.
    Strong scaling: the job does not grow with the GPU count.  Only rank 0 calls; the other ranks wait at a CPU barrier."""
    torch, K, R = cx.torch, cx.K, cx.R
    nd = cx.world
    mask = (1 << nd) - 1
    rec = None
    if cx.rank == 0:
        n_distinct = 8
        faces = [cx.pinned(R.img_mix(2048, 2048, 100 + i)).numpy() for i in range(n_distinct)]   # 8 distinct faces stand for the 384 (img_mix takes 0.3 s each)
        nb = 512 * 512
        out = np.empty((n_images * nb, 16), np.uint8)
        cx.torch.cuda.cudart().cudaHostRegister(out.ctypes.data, out.nbytes, 0)
        arr = (K.api.Image * n_images)()
        optr = (ctypes.c_void_p * n_images)()
        for i in range(n_images):
            f = faces[i % n_distinct]
            arr[i] = K.api.Image(f.ctypes.data, 2048, 2048, 0)
            optr[i] = out.ctypes.data + i * nb * 16
        params = K.api.make_params(LEVEL, device_mask=mask)
        aa = ctypes.c_uint32(0)
        K.api._check(K.lib().b200_uastc_encode(arr, min(n_images, 8 * nd), ctypes.byref(params), optr, ctypes.byref(aa)))   # warm-up: contexts, lanes, staging
        l0 = K.launch_count()
        t0 = time.perf_counter()
        K.api._check(K.lib().b200_uastc_encode(arr, n_images, ctypes.byref(params), optr, ctypes.byref(aa)))
        dt = time.perf_counter() - t0
        launches = K.launch_count() - l0
        o = out.reshape(n_images, nb, 16)
        same = all((o[i] == o[i % n_distinct]).all() for i in range(n_distinct, n_images))
        pix = n_images * 2048 * 2048
        rec = {"workload": C4_NAME if n_images == 384 else "C4 scaled to %d faces" % n_images, "metric": "uastc_encode_Mpix_per_s", "unit": "Mpix/s", "scaling": "strong",
               "value": pix / dt / 1e6, "ms_per_step": dt * 1e3, "steps": 1, "gpus_active": nd, "device_mask": mask,
               "note": "value == e2e: one b200_uastc_encode call, host buffers (8 distinct pinned faces referenced 48 times each; output gathered into one registered 1.6 GB host buffer), "
                       "wall clock around the call",
               "e2e": {"value": pix / dt / 1e6, "unit": "Mpix/s", "h2d_bytes_per_step": n_images * 2048 * 2048 * 4, "d2h_bytes_per_step": n_images * nb * 16,
                       "api": "b200_uastc_encode(device_mask=0x%x)" % mask},
               "gpu_launches": launches, "repeated_faces_identical": bool(same)}
        if R.have_ref():
            cores = os.cpu_count() or 1
            px = R.ref_extract_blocks(faces[0])
            t0 = time.perf_counter()
            want = R.ref_encode_blocks(px, LEVEL, cores)
            dtr = time.perf_counter() - t0
            rec["parity_vs_reference"] = {"encode_bit_exact": bool((o[0] == want).all() and (o[n_images - n_distinct] == want).all()), "blocks_checked": 2 * nb,
                                          "sample": "face 0 and face %d (same content) against the reference" % (n_images - n_distinct)}
            rec["cpu_baseline"] = {"value": 16 * nb / dtr / 1e6, "unit": "Mpix/s", "cores": cores, "kind": "reference",
                                   "sample": "face 0 of 384 (262144 blocks): reference encode_uastc level 2 over %d threads (nproc = %d); faces are independent" % (cores, cores)}
        # ---- supercompression of finished faces under the encode of later ones (b200_uastc_encode_stream).  One device,
        # 48 faces, Zstd level 1 through pyarrow on the calling thread: serial = encode, then compress every face; streamed = compress
        # face i inside its completion callback while faces i+1.. are on the GPU.
        try:
            import pyarrow as pa
            codec = pa.Codec("zstd", compression_level=1)
            n_z = min(48, n_images)
            p1 = K.api.make_params(LEVEL)
            sizes = [0] * n_z

            def compress(i):
                sizes[i] = len(codec.compress(memoryview(out[i * nb:(i + 1) * nb].reshape(-1)), asbytes=True))

            t0 = time.perf_counter()
            K.api._check(K.lib().b200_uastc_encode(arr, n_z, ctypes.byref(p1), optr, ctypes.byref(aa)))
            t1 = time.perf_counter()
            for i in range(n_z):
                compress(i)
            t2 = time.perf_counter()
            serial_sizes = list(sizes)
            cb = K.api.IMAGE_DONE_FN(lambda i, _u: compress(int(i)))
            K.api._check(K.lib().b200_uastc_encode_stream(arr, n_z, None, ctypes.byref(p1), optr, ctypes.byref(aa), cb, None))
            t3 = time.perf_counter()
            rec["zstd_under_encode"] = {"faces": n_z, "codec": "zstd level 1 (pyarrow), one host thread", "encode_ms": (t1 - t0) * 1e3, "zstd_ms": (t2 - t1) * 1e3,
                                        "serial_ms": (t2 - t0) * 1e3, "streamed_ms": (t3 - t2) * 1e3, "same_compressed_sizes": serial_sizes == sizes,
                                        "api": "b200_uastc_encode_stream: a completion callback per face"}
        except ImportError:
            pass
        cx.torch.cuda.cudart().cudaHostUnregister(out.ctypes.data)
    cx.cpu_barrier()
    return rec


def bench_c5(cx, args, with_cpu_targets=False):
    """HBM-streaming roofline run: 2^26 device-resident UASTC blocks -> BC7 (and the other block targets on a 2^24-block slice)."""
    torch, K, R = cx.torch, cx.K, cx.R
    n = args.blocks
    base, desc = make_uastc_corpus()
    g = torch.Generator(device="cpu").manual_seed(1000 + cx.rank)
    d_base = torch.from_numpy(base).to(cx.dev)
    idx = torch.randint(0, d_base.shape[0], (n,), generator=g).to(cx.dev)
    d_src = d_base[idx].contiguous()          # [n,16] u8, resident in HBM; 1 GiB >> 126 MB L2
    del idx
    d_dst = torch.empty_like(d_src)
    d_bad = torch.zeros(1, dtype=torch.int64, device=cx.dev)
    st = cx.st

    def step(ev):
        if ev: ev["bc7"][0].record()
        K.transcode_dev(K.TTF_BC7_RGBA, d_src.data_ptr(), d_dst.data_ptr(), n, st, d_bad.data_ptr())
        if ev: ev["bc7"][1].record()

    ms_per_step, per, launches, clocks = timed_loop(cx, step, max(args.steps, 10), max(3, args.warmup), ("bc7",))
    assert d_bad.item() == -1, "transcode flagged an invalid block"
    value = 32.0 * n * cx.world / (ms_per_step * 1e-3) / 1e9
    parity = None
    if R.have_ref():
        sub = slice(0, 1 << 18)
        parity = {"bc7_bit_exact": bool((R.ref_transcode(d_src[sub].cpu().numpy(), 0)[0] == d_dst[sub].cpu().numpy()).all()), "blocks_checked": 1 << 18,
                  "full_size_sweep": "profiles/r7_parity_sweep.json (2^26 blocks to BC7 and to ASTC, 0 mismatches)"}
    ref_head = d_dst[:1 << 14].cpu().numpy().copy()

    # e2e: same metric through the C-ABI host call, pinned host buffers, H2D + D2H inside the timed region
    ne = args.e2e_blocks
    h_src = torch.empty((ne, 16), dtype=torch.uint8).pin_memory()
    h_src.copy_(d_src[:ne].cpu())
    h_dst = torch.empty((ne, 16), dtype=torch.uint8).pin_memory()
    spans = [(0, 0, ne)]
    hs, hd = h_src.numpy(), h_dst.numpy().reshape(-1)
    e2e_s = e2e_loop(cx, lambda: K.transcode(hs, K.TTF_BC7_RGBA, spans=spans, dst=hd), args.steps)
    e2e_val = 32.0 * ne * cx.world / e2e_s / 1e9
    e2e_ok = bool((h_dst[:1 << 14].numpy() == ref_head).all())

    # The other block targets on a 2^24-block slice of the same input, device-resident.  bytes = 16 read + 8 or 16 written.
    targets = {}
    if cx.rank == 0:
        nt = min(n, 1 << 24)
        names = [("astc", K.TTF_ASTC_4x4_RGBA), ("bc1", K.TTF_BC1_RGB), ("bc3", K.TTF_BC3_RGBA), ("bc4", K.TTF_BC4_R), ("bc5", K.TTF_BC5_RG),
                 ("etc1", K.TTF_ETC1_RGB), ("etc2", K.TTF_ETC2_RGBA), ("eac_r11", K.TTF_ETC2_EAC_R11), ("eac_rg11", K.TTF_ETC2_EAC_RG11)]
        for name, tf in names:
            ob = K.target_block_bytes(tf)
            nn = n if tf == K.TTF_ASTC_4x4_RGBA else nt
            for _ in range(2):
                K.transcode_dev_ex(tf, 0, True, d_src.data_ptr(), d_dst.data_ptr(), nn, st, d_bad.data_ptr())
            a, b = cx.ev(), cx.ev()
            reps = 5
            a.record()
            for _ in range(reps):
                K.transcode_dev_ex(tf, 0, True, d_src.data_ptr(), d_dst.data_ptr(), nn, st, d_bad.data_ptr())
            b.record(); torch.cuda.synchronize()
            ms = a.elapsed_time(b) / reps
            ent = {"gbs": (16.0 + ob) * nn / ms / 1e6, "ms": ms, "blocks": nn, "bytes_per_block": 16 + ob,
                   "frac_of_hbm_peak": (16.0 + ob) * nn / ms / 1e6 / cx.pk["hbm_gbs"]}
            if R.have_ref():
                ns = 1 << 16
                hb = np.ascontiguousarray(d_src[:ns].cpu().numpy())
                got = d_dst.reshape(-1)[:ns * ob].cpu().numpy().reshape(ns, ob)
                want = R.ref_transcode(hb, 1)[0] if tf == K.TTF_ASTC_4x4_RGBA else R.ref_transcode_hinted(hb, tf)[0]
                ent["parity_vs_reference"] = bool((got == want).all())
                if with_cpu_targets and tf != K.TTF_ASTC_4x4_RGBA:
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
    if cx.rank != 0:
        return None
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
               "sample": "%d blocks -> BC7, unmodified reference transcoder over %d threads (nproc = %d), %d reps" % (nb, cores, cores, reps)}
    rl = cx.hbm_roofline("k_transcode_tiled<0>@c5", "k_transcode_tiled<BC7>", n, per["bc7"], 32.0 * n)
    return {"metric": "uastc_to_bc7_transcode_GBps", "value": value, "unit": "GB/s", "n_gpus": cx.world, "steps": max(args.steps, 10),
            "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": C5_NAME, "blocks_per_gpu": n, "corpus": desc, "l2": "inputs (1 GiB) larger than L2, no flush needed",
                       "parity_vs_reference": parity, "build": cx.build},
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": "GB/s", "h2d_bytes_per_step": 16 * ne, "d2h_bytes_per_step": 16 * ne,
                    "blocks": ne, "api": "b200_uastc_transcode (host pinned buffers)", "matches_device_path": e2e_ok},
            "gpu_launches": launches, "roofline": rl, "other_targets": targets, "cpu_baseline": cpu}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--size", type=int, default=8192, help="c2: base level edge (8192 = the named config)")
    ap.add_argument("--blocks", type=int, default=1 << 26)
    ap.add_argument("--e2e-blocks", type=int, default=1 << 25)
    ap.add_argument("--c4-images", type=int, default=384)
    ap.add_argument("--no-configs", action="store_true", help="c2: skip the records of the other named configs")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    cx = Ctx(args)
    wl = args.workload
    line = None
    if wl == "c2":
        line = bench_c2(cx, args)
        if not args.no_configs:
            cfgs = {}
            cx.cpu_barrier()
            if cx.rank == 0 and cx.world == 1:   # single-GPU runs carry every named config; multi-GPU runs only the sharded job
                cfgs["c1"] = bench_c1(cx, args)
                cfgs["c3_l3"], cfgs["c3_l4_rdo"] = bench_c3(cx, args)
                c5 = bench_c5(cx, args)
                cfgs["c5"] = {k: c5[k] for k in ("metric", "value", "unit", "ms_per_step", "steps", "config", "e2e", "gpu_launches", "roofline", "other_targets", "cpu_baseline", "clocks")}
            c4 = bench_c4(cx, args, args.c4_images)
            if cx.rank == 0:
                cfgs["c4"] = c4
                line["configs"] = cfgs
                line["sharded_job"] = {k: c4[k] for k in ("workload", "scaling", "value", "unit", "ms_per_step", "gpus_active", "device_mask", "parity_vs_reference") if k in c4}
    elif wl == "c5":
        line = bench_c5(cx, args, with_cpu_targets=True)
    else:
        if wl == "c4":
            rec = bench_c4(cx, args, args.c4_images)
        elif wl == "c1":
            rec = bench_c1(cx, args) if cx.rank == 0 else None
        else:
            rec = None
            if cx.rank == 0:
                l3, rdo = bench_c3(cx, args)
                rec = dict(l3)
                rec["level4_rdo"] = rdo
        cx.cpu_barrier()
        if cx.rank == 0:
            line = {"metric": rec["metric"], "value": rec["value"], "unit": rec["unit"], "n_gpus": cx.world, "steps": rec.get("steps", args.steps), "warmup": 3,
                    "ms_per_step": rec.get("ms_per_step", rec.get("kernel_ms")), "higher_is_better": True, "scaling": rec.get("scaling", "weak"), "vs_baseline": None,
                    "dtype": "u8", "data": "synthetic", "config": {"workload": rec["workload"], "build": cx.build, "parity_vs_reference": rec.get("parity_vs_reference")}}
            for k in ("e2e", "gpu_launches", "roofline", "cpu_baseline", "clocks", "level4_rdo", "gpus_active", "device_mask", "note", "zstd_under_encode"):
                if k in rec:
                    line[k] = rec[k]
    if cx.rank == 0 and line is not None:
        print(json.dumps(line))
    if cx.world > 1:
        cx.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
