#!/bin/bash
# Development aid: the encode probe over the library and every variant under ktx_software_b200/variants/, twice, on one box.
for rep in 1 2; do
for v in "" $(ls ktx_software_b200/variants/ 2>/dev/null | sed -n 's/^libuastc_\(.*\)\.so$/\1/p'); do
  if [ -n "$v" ]; then export B200_UASTC_LIB=$PWD/ktx_software_b200/variants/libuastc_$v.so; else unset B200_UASTC_LIB; fi
  echo "== variant ${v:-HEAD}"
  PROBE_KINDS=mix python tools/enc_probe.py 4096 ${AB_LEVELS:-1,2,3} 2>&1 | tail -3
  PROBE_KINDS=noise,smooth python tools/enc_probe.py 2048 2 2>&1 | tail -2
  [ -n "$AB_BIG" ] && PROBE_KINDS=mix python tools/enc_probe.py 8192 2 2>&1 | tail -1
done
done
