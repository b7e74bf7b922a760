NCU="ncu --set full --clock-control none --import-source on"
PROBE_KINDS=mix python tools/enc_probe.py 1024 4 > gpurun_out/l4_plain.log 2>&1 &&
PROBE_KINDS=mix $NCU -k regex:k_encode_jobs -s 1 -c 1 -o gpurun_out/l4 python tools/enc_probe.py 1024 4 > gpurun_out/l4_ncu.log 2>&1
python tools/ncu_summary.py gpurun_out/l4.ncu-rep > gpurun_out/l4_summary.txt 2>&1
python tools/ncu_by_func.py gpurun_out/l4.ncu-rep ktx_software_b200/libuastc_b200.so k_encode >> gpurun_out/l4_summary.txt 2>/dev/null
rm -f gpurun_out/*.ncu-rep
