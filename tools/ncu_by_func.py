#!/usr/bin/env python3
"""Development aid: attribute an ncu report's per-SASS-instruction counters to the device functions inside the kernel.
usage: ncu_by_func.py report.ncu-rep lib.so [kernel-substring]
Uses `ncu --page source --print-source sass --csv` for the counters and the cubin's symbol table for function ranges."""
import csv, io, os, re, subprocess, sys, tempfile

rep, so = sys.argv[1], sys.argv[2]
want = sys.argv[3] if len(sys.argv) > 3 else "k_encode"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
syms = subprocess.run(["readelf", "-sW", cubin], capture_output=True, text=True).stdout
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
# split per kernel
chunks = re.split(r'(?m)^"Kernel Name",', out)
for ch in chunks[1:]:
    name = ch.split("\n", 1)[0]
    if want not in name:
        continue
    rows = list(csv.reader(io.StringIO(ch.split("\n", 1)[1])))
    hdr = rows[0]
    ix = {h: i for i, h in enumerate(hdr)}
    base = int(rows[1][0], 16)
    # kernel mangled name: find FUNC symbols "$<kernel>$<func>" whose kernel matches template args in `name`
    m = re.search(r"(k_\w+)<((?:\((?:int|bool)\)\d+(?:, )?)+)>", name)
    kfilter = (m.group(1) + "I" + "".join("L%s%sE" % ("b" if t == "bool" else "i", v) for t, v in re.findall(r"\((int|bool)\)(\d+)", m.group(2)))) if m else want
    funcs = []
    for line in syms.splitlines():
        p = line.split()
        if len(p) >= 8 and p[3] == "FUNC" and kfilter in p[-1]:
            size = int(p[2], 0)
            nm = p[-1].split("$")[-1] if p[-1].startswith("$") else "<kernel body>"
            funcs.append((int(p[1], 16), size, nm, p[-1].startswith("$")))
    ksym = [f for f in funcs if not f[3]][0]
    subs = sorted([f for f in funcs if f[3]])
    agg = {}
    tot = [0, 0, 0, 0]
    for r in rows[1:]:
        if len(r) < len(hdr):
            continue
        off = int(r[0], 16) - base + ksym[0]
        fn = "<kernel body>"
        for o, s, n, _ in subs:
            if o <= off < o + s:
                fn = n
                break
        a = agg.setdefault(fn, [0, 0, 0, 0])
        vals = [int(r[ix["Instructions Executed"]] or 0), int(r[ix["Thread Instructions Executed"]] or 0),
                int(r[ix["# Samples"]] or 0), int(r[ix["stall_no_inst"]] or 0)]
        for k in range(4):
            a[k] += vals[k]; tot[k] += vals[k]
    print(name)
    print("%-28s %8s %7s %7s %7s %9s" % ("function", "inst%", "thr/in", "samp%", "noinst%", "size"))
    sz = {n: s for _, s, n, _ in subs}
    for fn, a in sorted(agg.items(), key=lambda kv: -kv[1][2]):
        d = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip().split("(")[0]
        print("%-28s %8.2f %7.1f %7.2f %7.1f %9s" % (d[-28:], 100.0 * a[0] / max(1, tot[0]), a[1] / max(1, a[0]),
                                                  100.0 * a[2] / max(1, tot[2]), 100.0 * a[3] / max(1, a[2]), sz.get(fn, "")))
    print("total warp-inst %d, samples %d, no_inst share %.1f%%" % (tot[0], tot[2], 100.0 * tot[3] / max(1, tot[2])))
