#!/bin/bash
# Development aid: A/B of two builds of the library on the same box.  usage: tools/ab_c5.sh <other.so>
python -m pytest tests/test_transcode_gpu.py tests/test_decode_gpu.py -x -q -m gpu 2>&1 | tail -1
for i in 1 2; do
  python bench.py --workload c5 2>/dev/null | grep -o '"achieved": [0-9.]*' | head -1
  B200_UASTC_LIB=$1 python bench.py --workload c5 2>/dev/null | grep -o '"achieved": [0-9.]*' | head -1 | sed 's/^/other /'
done
python tools/decode_probe.py 16384 13
B200_UASTC_LIB=$1 python tools/decode_probe.py 16384 13 | sed 's/^/other /'
