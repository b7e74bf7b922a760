#!/usr/bin/env python3
"""Parity at config size: the CUDA path through the C ABI against the UNMODIFIED reference (oracle/_ref) on all host cores.

  python tools/parity_sweep.py [--quick] [--out gpurun_out/r4_parity_sweep.json]

Runs on the GPU box (needs a B200 and oracle/_ref/libref_uastc.so; never reads /root/reference).  For every BASELINE.json
config it encodes / transcodes the named synthetic input with both implementations and compares ALL blocks:

  C1   img_mix / img_noise / img_smooth (2048,2048,1), level 2                        -> bit-exact
  C2   img_mix(8192,8192,2) + 13 mips, level 2; every block -> BC7 and -> ASTC        -> bit-exact
  C3   img_mix(4096,4096,3): level 3 no-RDO (bit-exact tier);
       level 4 + RDO lambda 1 (dict 4096, 4 jobs, FavorSimplerModes as libktx sets it) -> PSNR of both, |d| <= 0.01 dB,
                                                                                          per-block mode agreement, zstd-19 sizes
  C5   16 x img_mix(1024,1024,200+j) at level 3 (all 19 modes present), expanded by seeded block permutations to 2^26 blocks,
       -> BC7 and -> ASTC                                                             -> bit-exact
  L0/L1/L3 sweeps: img_mix(4096,4096,7) + img_noise(1280,1280,5) at levels 0 and 1; img_noise at level 3.

Test infrastructure (the oracle is only the checker here).  --quick runs a bounded slice (tests/test_parity_sweep_gpu.py).
"""
import argparse, json, os, sys, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def psnr(a, b):
    mse = ((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean()
    return 99.0 if mse == 0 else float(10 * np.log10(255.0 ** 2 / mse))


def zstd_size(buf, level=19):
    import pyarrow as pa
    return int(pa.Codec("zstd", compression_level=level).compress(memoryview(np.ascontiguousarray(buf).reshape(-1)), asbytes=True).__len__())


def mip_chain(base):
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


def mode_hist(blocks):
    import refutil as R
    return np.bincount(R.block_modes(blocks), minlength=20)[:20].tolist()


class Sweep:
    def __init__(self, quick):
        import ktx_software_b200 as K
        import refutil as R
        self.K, self.R, self.quick = K, R, quick
        K.init()
        assert R.have_ref(), "oracle/_ref/libref_uastc.so is needed (the unmodified reference is the checker)"
        self.cores = os.cpu_count() or 1
        self.rows = []

    def log(self, row):
        self.rows.append(row)
        print(json.dumps(row), flush=True)

    def encode_pair(self, imgs, flags):
        """-> (ours [n,16], reference [n,16], pixels [n,64], seconds ours, seconds reference)"""
        K, R = self.K, self.R
        t0 = time.perf_counter()
        outs, _ = K.encode_images(imgs, flags)
        t1 = time.perf_counter()
        px = np.concatenate([R.ref_extract_blocks(im) for im in imgs])
        t2 = time.perf_counter()
        want = R.ref_encode_blocks(px, flags, self.cores)
        t3 = time.perf_counter()
        return np.concatenate(outs), want, px, t1 - t0, t3 - t2

    def check_encode(self, name, imgs, flags, need_all_modes=False):
        got, want, px, t_ours, t_ref = self.encode_pair(imgs, flags)
        bad = np.nonzero((got != want).any(1))[0]
        row = {"config": name, "what": "encode", "flags": flags, "blocks": int(len(px)), "mismatches": int(len(bad)),
               "first_bad": bad[:8].tolist(), "modes": mode_hist(want), "seconds_ours_host_api": round(t_ours, 3),
               "seconds_reference": round(t_ref, 3), "reference_threads": self.cores}
        if len(bad):
            row["psnr_ours"] = psnr(self.R.ref_unpack_rgba(got)[0], px)
            row["psnr_reference"] = psnr(self.R.ref_unpack_rgba(want)[0], px)
        if need_all_modes:
            row["all_19_modes_present"] = bool(all(c > 0 for c in row["modes"][:19]))
        self.log(row)
        return got, want, px

    def check_transcode(self, name, blocks):
        K, R = self.K, self.R
        blocks = np.ascontiguousarray(blocks).reshape(-1, 16)
        n = len(blocks)
        for fmt, tf, label in ((0, K.TTF_BC7_RGBA, "bc7"), (1, K.TTF_ASTC_4x4_RGBA, "astc")):
            t0 = time.perf_counter()
            got = K.transcode(blocks, tf).reshape(-1, 16)
            t1 = time.perf_counter()
            want = np.empty_like(blocks)
            rc = R.ref().ref_transcode_mt(blocks.ctypes.data_as(R._u8p), n, fmt, want.ctypes.data_as(R._u8p), self.cores)
            t2 = time.perf_counter()
            assert rc == -1, "the reference rejected block %d" % rc
            nbad = 0
            first = []
            step = 1 << 22
            for o in range(0, n, step):   # chunked compare: bounded temporaries at 2^26 blocks
                b = np.nonzero((got[o:o + step] != want[o:o + step]).any(1))[0]
                nbad += len(b)
                if len(b) and not first:
                    first = (b[:8] + o).tolist()
            self.log({"config": name, "what": "transcode_" + label, "blocks": int(n), "mismatches": int(nbad), "first_bad": first,
                      "seconds_ours_host_api": round(t1 - t0, 3), "seconds_reference": round(t2 - t1, 3), "reference_threads": self.cores})

    # ------------------------------------------------------------------------------------------------------
    def c1(self):
        R = self.R
        s = 512 if self.quick else 2048
        for fam, f in (("img_mix", R.img_mix), ("img_noise", R.img_noise), ("img_smooth", R.img_smooth)):
            self.check_encode("C1 %s(%d,%d,1) level 2" % (fam, s, s), [f(s, s, 1)], 2)

    def c2(self):
        R = self.R
        s = 1024 if self.quick else 8192
        imgs = mip_chain(R.img_mix(s, s, 2))
        got, want, _ = self.check_encode("C2 img_mix(%d,%d,2) + %d mips, level 2" % (s, s, len(imgs) - 1), imgs, 2)
        self.check_transcode("C2 all levels", want)

    def c3(self):
        K, R = self.K, self.R
        s = 512 if self.quick else 4096
        img = R.img_mix(s, s, 3)
        self.check_encode("C3 img_mix(%d,%d,3) level 3 no RDO" % (s, s), [img], 3)
        # level 4 without RDO (PSNR tier; reported bit-exact where it is)
        got4, want4, px = self.check_encode("C3 img_mix(%d,%d,3) level 4 no RDO" % (s, s), [img], 4)
        # level 4 + RDO as libktx configures it: lambda 1, dict 4096, jobs = min(4, threadCount), FavorSimplerModes (512) set
        # because RDO is on (lib/src/basis_encode.cpp:900-903)
        flags = 4 | 512
        t0 = time.perf_counter()
        outs, _ = K.encode_images([img], flags, rdo=True, rdo_lambda=1.0, total_jobs=4)
        t1 = time.perf_counter()
        ours = outs[0]
        ref_pre = R.ref_encode_blocks(px, flags, self.cores)
        t2 = time.perf_counter()
        ref = R.ref_rdo(ref_pre, px, flags, 4, 4, lam=1.0, dict_size=4096)
        t3 = time.perf_counter()
        pa, pb = psnr(R.ref_unpack_rgba(ours)[0], px), psnr(R.ref_unpack_rgba(ref)[0], px)
        p_pre = psnr(R.ref_unpack_rgba(ref_pre)[0], px)
        mo, mr = R.block_modes(ours), R.block_modes(ref)
        bad = np.nonzero((ours != ref).any(1))[0]
        self.log({"config": "C3 img_mix(%d,%d,3) level 4 + RDO lambda 1.0 (dict 4096, 4 jobs, flags 4|512)" % (s, s), "what": "encode+rdo",
                  "blocks": int(len(px)), "mismatches": int(len(bad)), "first_bad": bad[:8].tolist(),
                  "psnr_ours_db": pa, "psnr_reference_db": pb, "psnr_delta_db": abs(pa - pb), "psnr_within_0.01_db": bool(abs(pa - pb) <= 0.01),
                  "psnr_reference_before_rdo_db": p_pre, "mode_agreement": float((mo == mr).mean()),
                  "blocks_modified_by_rdo_reference": int((ref != ref_pre).any(1).sum()),
                  "zstd19_bytes_ours": zstd_size(ours), "zstd19_bytes_reference": zstd_size(ref), "zstd19_bytes_reference_before_rdo": zstd_size(ref_pre),
                  "seconds_ours_host_api": round(t1 - t0, 3), "seconds_reference_encode": round(t2 - t1, 3),
                  "seconds_reference_rdo_4_jobs": round(t3 - t2, 3), "reference_threads": self.cores})

    def c5(self):
        R = self.R
        nb, s = (2, 256) if self.quick else (16, 1024)
        imgs = [R.img_mix(s, s, 200 + j) for j in range(nb)]
        got, want, _ = self.check_encode("C5 base textures: %d x img_mix(%d,%d,200+j) level 3" % (nb, s, s), imgs, 3, need_all_modes=True)
        n = (1 << 18) if self.quick else (1 << 26)
        per = len(want)
        reps = n // per
        rng = np.random.default_rng(5)
        big = np.empty((n, 16), np.uint8)
        for k in range(reps):   # seeded block permutations of the base set (transcode is per block: validity is preserved)
            big[k * per:(k + 1) * per] = want[rng.permutation(per)]
        self.check_transcode("C5 %d blocks (%d permutations of the %d base blocks)" % (n, reps, per), big)

    def levels(self):
        R = self.R
        s, sn = (512, 256) if self.quick else (4096, 1280)
        mix, noise = R.img_mix(s, s, 7), R.img_noise(sn, sn, 5)
        for lvl in (0, 1):
            self.check_encode("level %d img_mix(%d,%d,7)" % (lvl, s, s), [mix], lvl)
            self.check_encode("level %d img_noise(%d,%d,5)" % (lvl, sn, sn), [noise], lvl)
        self.check_encode("level 3 img_noise(%d,%d,5)" % (sn, sn), [noise], 3)
        self.check_encode("level 2 adversarial families x2048", [np.ascontiguousarray(
            R.adversarial_blocks(11, 256 if self.quick else 2048).reshape(-1, 4, 4, 4).transpose(1, 0, 2, 3).reshape(4, -1, 4))], 2)

    def run(self, which):
        for name in which:
            getattr(self, name)()
        enc = [r for r in self.rows if r["what"].startswith("encode")]
        xc = [r for r in self.rows if r["what"].startswith("transcode")]
        bit_exact_rows = [r for r in self.rows if r["what"] != "encode+rdo" and "level 4" not in r["config"]]
        return {"quick": self.quick, "host_cores": self.cores,
                "encode_blocks_checked": int(sum(r["blocks"] for r in enc)), "transcode_blocks_checked": int(sum(r["blocks"] for r in xc)),
                "bit_exact_tier_mismatches": int(sum(r["mismatches"] for r in bit_exact_rows)),
                "all_mismatches": int(sum(r["mismatches"] for r in self.rows)), "rows": self.rows}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--only", default="c1,c2,c3,c5,levels")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r4_parity_sweep.json"))
    a = ap.parse_args()
    res = Sweep(a.quick).run(a.only.split(","))
    import ktx_software_b200 as K
    res["build"] = K.build_id()   # the library the sweep ran through (SHA-1 of its sources, as in bench.py's records)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(res, open(a.out, "w"), indent=1)
    print("SUMMARY", json.dumps({k: v for k, v in res.items() if k != "rows"}))
    return 0 if res["bit_exact_tier_mismatches"] == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
