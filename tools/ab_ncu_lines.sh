#!/bin/bash
# Development aid, ON THE GPU BOX: one ncu capture of k_encode_jobs (4096^2 img_mix, level $2), by device function and by source line of function $1.
F=$1; L=${2:-2}
NCU="ncu --set full --clock-control none --import-source on"
PROBE_KINDS=mix python tools/enc_probe.py 4096 $L > gpurun_out/ab_plain.log 2>&1 &&
PROBE_KINDS=mix $NCU -k regex:k_encode_jobs -s 1 -c 1 -o gpurun_out/ab_head python tools/enc_probe.py 4096 $L > gpurun_out/ab_ncu_head.log 2>&1
python tools/ncu_summary.py gpurun_out/ab_head.ncu-rep > gpurun_out/ab_head_summary.txt 2>&1
python tools/ncu_by_func.py gpurun_out/ab_head.ncu-rep ktx_software_b200/libuastc_b200.so k_encode >> gpurun_out/ab_head_summary.txt 2>/dev/null
python tools/ncu_by_line.py gpurun_out/ab_head.ncu-rep ktx_software_b200/libuastc_b200.so k_encode_jobs $F 60 > gpurun_out/ab_head_lines.txt 2>&1
rm -f gpurun_out/*.ncu-rep
