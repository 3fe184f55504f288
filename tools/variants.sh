#!/bin/bash
# Build-time variants of the kernels (-D flags) timed side by side at C3 on the GPU box.
#   build:  bash tools/variants.sh build "name1:-DFLAG=1 -DOTHER=2" "name2:..."     (CPU box; output tools/variants/lib_<name>.so)
#   run:    bash tools/variants.sh run [bench args]                                   (under gpurun; output gpurun_out/variants.log)
set -e
cd "$(dirname "$0")/.."
if [ "$1" = build ]; then
  shift
  mkdir -p tools/variants
  rm -f tools/variants/lib_*.so
  for spec in "$@"; do
    name="${spec%%:*}"; flags="${spec#*:}"
    ( nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -Xcompiler -fPIC -ccbin /usr/bin/g++ $flags \
        -c ecolab_b200/csrc/engine.cu -o tools/variants/engine_$name.o &&
      nvcc -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ -shared -o tools/variants/lib_$name.so tools/variants/engine_$name.o ecolab_b200/csrc/model.o -cudart static &&
      rm tools/variants/engine_$name.o ) &
  done
  wait
  ls -la tools/variants
else
  shift
  B="python bench.py --steps 10 --warmup 3 ${VARGRID:---nx 1024 --ny 1024} --no-cpu --no-e2e --no-extra --no-verify $*"
  : > gpurun_out/variants.log
  for lib in default tools/variants/lib_*.so; do
    if [ $lib = default ]; then unset ECOLAB_B200_LIB; else export ECOLAB_B200_LIB=$PWD/$lib; fi
    echo "== $lib" >> gpurun_out/variants.log
    timeout 200 $B 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'): j=json.loads(l); print(j['roofline']['kernels_ms_per_step'], 'value %.3e' % j['value'])
    else: print(l.rstrip()[:160])" >> gpurun_out/variants.log
  done
fi
