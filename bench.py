#!/usr/bin/env python3
"""bench.py -- the driver's measurement contract for the UASTC hot path (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload transcode|encode]

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path over one batch of synthetic input.
"""
import argparse, json, os, subprocess, sys, threading, time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


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


def make_uastc_corpus(n_base_images=4, size=512, level=2):
    """Reference-encoded UASTC blocks of img_mix textures (SURVEY 8(d) C5 recipe, scaled): realistic mode mix."""
    import refutil as R
    if R.have_ref():
        blocks = [R.ref_encode_blocks(R.ref_extract_blocks(R.img_mix(size, size, 200 + j)), level, os.cpu_count() or 8)
                  for j in range(n_base_images)]
        return np.concatenate(blocks), "img_mix %dx%d x%d encoded by the reference at level %d" % (size, size, n_base_images, level)
    g = np.load(os.path.join(ROOT, "tests", "golden", "transcode_golden.npz"))
    return g["uastc"], "tests/golden/transcode_golden.npz corpus (oracle/_ref absent)"


def run_reference(args):
    """--impl reference: the reference's own CPU transcoder (oracle/_ref, unmodified sources), all host threads."""
    import refutil as R
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    if not R.have_ref():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_uastc.so not present"}))
        return 0
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
            "config": {"workload": WORKLOAD_NAME, "sample": "%d blocks per step" % n, "corpus": desc},
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "reference",
                             "sample": "%d UASTC blocks -> BC7 per step, ref_transcode_mt over %d threads" % (n, cores)},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


WORKLOAD_NAME = "C5: batch transcode of 1024 UASTC 1024x1024 textures to BC7 (2^26 blocks, 1 GiB in / 1 GiB out per GPU)"


def run_ours(args):
    import torch
    import torch.distributed as dist
    import ktx_software_b200 as K
    import refutil as R

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
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

    for _ in range(max(3, args.warmup)):
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
    total_ms = t.item()
    ms_per_step = total_ms / args.steps
    value = 32.0 * n * world / (ms_per_step * 1e-3) / 1e9

    # parity spot check of this very output against the reference (not timed)
    parity = None
    if R.have_ref():
        sub = slice(0, 1 << 16)
        parity = bool((R.ref_transcode(d_src[sub].cpu().numpy(), 0)[0] == d_dst[sub].cpu().numpy()).all())

    # ---- e2e: same metric through the C-ABI host call, pinned host buffers, H2D + D2H inside the timed region
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
                "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": WORKLOAD_NAME, "blocks_per_gpu": n, "corpus": desc,
                           "l2": "inputs (1 GiB) larger than L2, no flush needed", "parity_vs_reference": parity},
                "clocks": cs.summary(),
                "e2e": {"value": e2e_val, "unit": "GB/s", "h2d_bytes_per_step": 16 * ne, "d2h_bytes_per_step": 16 * ne,
                        "blocks": ne, "api": "b200_uastc_transcode (host pinned buffers)", "matches_device_path": e2e_ok},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                             "frac": achieved / pk["hbm_gbs"], "traffic": None, "peak_kind": pk_kind,
                             "kernel": "k_transcode<BC7>", "kernel_ms": kern_ms, "algorithmic_bytes_per_launch": 32 * n},
                "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--blocks", type=int, default=1 << 26)
    ap.add_argument("--e2e-blocks", type=int, default=1 << 25)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
