#!/bin/bash
# Run ON THE GPU BOX (through gpurun): plain run of each bench workload, then the ncu captures bench.py's roofline records quote
# (one kernel launch each, --set full).  The reports are reduced ON THE BOX to text summaries + gpurun_out/$1_ncu_counters.json
# (gpurun brings back at most 64 MiB); only the C2 encode report itself is kept.
P=gpurun_out/$1
NCU="ncu --set full --clock-control none --import-source on"
BUILD=$(python -c "import ktx_software_b200 as K; print(K.build_id())")
python bench.py --no-configs --steps 1 > ${P}_plain_c2.log 2>&1 &&
$NCU -k regex:k_encode_jobs -s 3 -c 1 -o ${P}_enc_c2 python bench.py --no-configs --steps 1 > ${P}_ncu_enc_c2.log 2>&1
python bench.py --no-configs --steps 1 > ${P}_plain_c2b.log 2>&1 &&
$NCU -k regex:k_transcode_tiled -s 6 -c 2 -o ${P}_xc_c2 python bench.py --no-configs --steps 1 > ${P}_ncu_xc_c2.log 2>&1
python bench.py --workload c5 --steps 3 > ${P}_plain_c5.log 2>&1 &&
$NCU -k regex:k_transcode_tiled -s 3 -c 1 -o ${P}_xc_c5 python bench.py --workload c5 --steps 3 > ${P}_ncu_xc_c5.log 2>&1
python bench.py --workload c1 --steps 2 > ${P}_plain_c1.log 2>&1 &&
$NCU -k regex:k_encode_jobs -s 3 -c 1 -o ${P}_enc_c1 python bench.py --workload c1 --steps 2 > ${P}_ncu_enc_c1.log 2>&1
PROBE_KINDS=mix python tools/enc_probe.py 4096 3 > ${P}_plain_l3.log 2>&1 &&
PROBE_KINDS=mix $NCU -k regex:k_encode_jobs -s 1 -c 1 -o ${P}_enc_c3l3 python tools/enc_probe.py 4096 3 > ${P}_ncu_enc_c3l3.log 2>&1
PROBE_KINDS=mix python tools/enc_probe.py 1024 4 > ${P}_plain_l4.log 2>&1 &&
PROBE_KINDS=mix $NCU -k regex:k_encode_blocks -s 1 -c 1 -o ${P}_enc_l4 python tools/enc_probe.py 1024 4 > ${P}_ncu_enc_l4.log 2>&1
for r in enc_c2 xc_c2 xc_c5 enc_c1 enc_c3l3 enc_l4; do
  [ -f ${P}_$r.ncu-rep ] || continue
  python tools/ncu_summary.py ${P}_$r.ncu-rep > ${P}_${r}_ncu_summary.txt 2>&1
  case $r in enc_*) python tools/ncu_by_func.py ${P}_$r.ncu-rep ktx_software_b200/libuastc_b200.so k_encode >> ${P}_${r}_ncu_summary.txt 2>/dev/null;; esac
done
python tools/ncu_counters.py $BUILD "k_encode_jobs@c2=${P}_enc_c2.ncu-rep:5592407" "k_transcode_tiled<0>@c2=${P}_xc_c2.ncu-rep:5592407" \
  "k_transcode_tiled<0>@c5=${P}_xc_c5.ncu-rep:67108864" "k_encode_jobs@c1=${P}_enc_c1.ncu-rep:262144" "k_encode_jobs@c3_l3=${P}_enc_c3l3.ncu-rep:1048576" \
  "k_encode_image<170>@c3_l4=${P}_enc_l4.ncu-rep:65536" > ${P}_counters.log 2>&1
cp profiles/ncu_counters.json ${P}_ncu_counters.json
for r in xc_c2 xc_c5 enc_c1 enc_c3l3 enc_l4; do rm -f ${P}_$r.ncu-rep; done
du -sh gpurun_out; tail -3 ${P}_counters.log | cut -c1-200
