#!/bin/bash
# Development aid: rebuild the in-tree CUDA library (so the snapshot carries a fresh .so), then run a command on the GPU box.
# usage: tools/gpu.sh [--gpus N] <timeout-seconds> '<command>'
set -e
cd "$(dirname "$0")/.."
G=""
if [ "$1" == "--gpus" ]; then G="--gpus $2"; shift 2; fi
python -m ktx_software_b200.build >/dev/null 2>/tmp/build_err.log || { cat /tmp/build_err.log | tail -30; exit 1; }
make -C oracle port >/dev/null
T=$1; shift
exec /usr/local/graft/bin/gpurun $G --timeout $T -- "$@"
