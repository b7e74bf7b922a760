#!/usr/bin/env python3
"""Development aid: attribute an ncu report's per-SASS counters of ONE device function to its CUDA source lines.
usage: ncu_by_line.py report.ncu-rep lib.so kernel-substring function-substring [top-n]
(needs the library built with -lineinfo; uses nvdisasm --print-line-info on the embedded cubin)"""
import csv, io, os, re, subprocess, sys, tempfile

rep, so, want, fwant = sys.argv[1:5]
topn = int(sys.argv[5]) if len(sys.argv) > 5 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
syms = subprocess.run(["readelf", "-sW", cubin], capture_output=True, text=True).stdout
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
for ch in re.split(r'(?m)^"Kernel Name",', out)[1:]:
    name = ch.split("\n", 1)[0]
    if want not in name:
        continue
    rows = list(csv.reader(io.StringIO(ch.split("\n", 1)[1])))
    hdr = rows[0]; ix = {h: i for i, h in enumerate(hdr)}
    base = int(rows[1][0], 16)
    m = re.search(r"(k_\w+)<((?:\((?:int|bool)\)\d+(?:, )?)+)>", name)
    kfilter = (m.group(1) + "I" + "".join("L%s%sE" % ("b" if t == "bool" else "i", v) for t, v in re.findall(r"\((int|bool)\)(\d+)", m.group(2)))) if m else want
    ksym = fsym = None
    for line in syms.splitlines():
        p = line.split()
        if len(p) >= 8 and p[3] == "FUNC" and kfilter in p[-1]:
            if not p[-1].startswith("$"): ksym = (int(p[1], 16), int(p[2], 0), p[-1])
            elif fwant in p[-1]: fsym = (int(p[1], 16), int(p[2], 0), p[-1])
    if fwant == 'KERNEL': fsym = ksym   # the kernel's own body (everything inlined)
    # line table of the function: walk nvdisasm's text for the kernel's section, track "//## File ..., line N"
    sec = dis.split(".section\t.text." + ksym[2], 1)[1].split(".section", 1)[0]
    cur = ("?", 0); lines = {}
    for l in sec.splitlines():
        mm = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if mm: cur = (os.path.basename(mm.group(1)), int(mm.group(2))); continue
        mm = re.search(r"/\*([0-9a-f]{4,})\*/\s+(\S.*?);", l)
        if mm: lines[int(mm.group(1), 16)] = (cur, mm.group(2))
    agg = {}; tot = [0, 0, 0]
    for r in rows[1:]:
        if len(r) < len(hdr): continue
        off = int(r[0], 16) - base           # offset inside the kernel's text section
        if not (fsym[0] - ksym[0] <= off < fsym[0] - ksym[0] + fsym[1]): continue
        key = lines.get(off, (("?", 0), ""))[0]
        a = agg.setdefault(key, [0, 0, 0, {}])
        v = [int(r[ix["Instructions Executed"]] or 0), int(r[ix["Thread Instructions Executed"]] or 0), int(r[ix["# Samples"]] or 0)]
        for k in range(3): a[k] += v[k]; tot[k] += v[k]
        op = r[1].split()[0] if r[1].split() else "?"
        if op.startswith("@"): op = r[1].split()[1]
        a[3][op.split(".")[0]] = a[3].get(op.split(".")[0], 0) + v[0]
    print(name, fsym[2][-40:])
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
        ops = ",".join("%s:%.0f%%" % (o, 100.0 * c / max(1, a[0])) for o, c in sorted(a[3].items(), key=lambda kv: -kv[1])[:4])
        print("%-22s:%-4d inst %5.2f%%  thr/in %4.1f  samp %5.2f%%  %s" % (key[0], key[1], 100.0 * a[0] / max(1, tot[0]), a[1] / max(1, a[0]), 100.0 * a[2] / max(1, tot[2]), ops))
