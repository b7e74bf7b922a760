#!/usr/bin/env python3
"""Turn ncu reports of the library's kernels into profiles/ncu_counters.json entries keyed by the library's build id, so that
bench.py quotes an instruction count or a DRAM traffic figure only for the build it actually loaded.

  python tools/ncu_counters.py <build-id> <key>=<report.ncu-rep>:<blocks> [...]

key = the name bench.py looks up ("k_encode_images<21>", "k_transcode_tiled<0>@c5", ...); blocks = 4x4 blocks the profiled launch
processed.  The first kernel of each report is used."""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, r = rows[0], rows[1], rows[2]
    d = {}
    for i, h in enumerate(hdr):
        d[h] = (r[i], units[i])
    return d


def num(d, k):
    v, u = d[k]
    v = float(v.replace(",", ""))
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "s": 1e3}.get(u, 1.0)
    return v * scale


def main():
    build = sys.argv[1]
    path = os.path.join(ROOT, "profiles", "ncu_counters.json")
    allc = json.load(open(path)) if os.path.exists(path) else {}
    ent = allc.setdefault(build, {})
    for spec in sys.argv[2:]:
        key, rest = spec.split("=", 1)
        rep, blocks = rest.rsplit(":", 1)
        d = raw(rep)
        blocks = int(blocks)
        wi = num(d, "smsp__inst_executed.sum")
        ent[key] = {"kernel": d["Kernel Name"][0], "blocks": blocks, "warp_inst": wi, "warp_inst_per_block": wi / blocks,
                    "dram_bytes": num(d, "dram__bytes_read.sum") + num(d, "dram__bytes_write.sum"),
                    "dram_bytes_read": num(d, "dram__bytes_read.sum"), "dram_bytes_write": num(d, "dram__bytes_write.sum"),
                    "issue_active_pct": num(d, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    "threads_per_inst": num(d, "smsp__thread_inst_executed_per_inst_executed.ratio"),
                    "ncu_duration_ms": num(d, "gpu__time_duration.sum"), "source": "profiles/%s (ncu --set full --clock-control none)" % os.path.basename(rep).replace(".ncu-rep", "_ncu_summary.txt")}
        print(key, json.dumps(ent[key]))
    json.dump(allc, open(path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
