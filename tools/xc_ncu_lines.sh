#!/bin/bash
# Development aid, ON THE GPU BOX: one ncu capture of k_transcode_tiled<BC7> on the C5 workload, summary + per-source-line attribution.
NCU="ncu --set full --clock-control none --import-source on"
python bench.py --workload c5 --steps 3 > gpurun_out/xc_plain.log 2>&1 &&
$NCU -k regex:k_transcode_tiled -s 3 -c 1 -o gpurun_out/xc_c5 python bench.py --workload c5 --steps 3 > gpurun_out/xc_ncu.log 2>&1
python tools/ncu_summary.py gpurun_out/xc_c5.ncu-rep > gpurun_out/xc_c5_summary.txt 2>&1
python tools/ncu_by_line.py gpurun_out/xc_c5.ncu-rep ktx_software_b200/libuastc_b200.so k_transcode_tiled KERNEL 70 > gpurun_out/xc_c5_lines.txt 2>&1
rm -f gpurun_out/*.ncu-rep
