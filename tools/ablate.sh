#!/bin/bash
# Cost attribution of generate_pipe_kernel: builds variants with one component removed (-DECO_ABLATE=bits:
# 1 no Philox, 2 no beta terms, 4 no ROUND, 8 no store) into tools/ablate/, and times each at C3 on the GPU box.
# build:  bash tools/ablate.sh build      (CPU box)      run:  bash tools/ablate.sh run   (under gpurun)
set -e
cd "$(dirname "$0")/.."
if [ "$1" = build ]; then
  mkdir -p tools/ablate
  for a in 1 2 4 8 15; do
    nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -Xcompiler -fPIC -ccbin /usr/bin/g++ -DECO_ABLATE=$a \
      -c ecolab_b200/csrc/engine.cu -o tools/ablate/engine_$a.o
    nvcc -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ -shared -o tools/ablate/lib_$a.so tools/ablate/engine_$a.o ecolab_b200/csrc/model.o -cudart static
    rm tools/ablate/engine_$a.o
  done
else
  B="python bench.py --steps 10 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e"
  : > gpurun_out/ablate.log
  for a in 0 1 2 4 8 15; do
    if [ $a = 0 ]; then unset ECOLAB_B200_LIB; else export ECOLAB_B200_LIB=$PWD/tools/ablate/lib_$a.so; fi
    echo "== ablate=$a" >> gpurun_out/ablate.log
    $B 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'): print(json.loads(l)['roofline']['kernels_ms_per_step'])
    else: print(l.rstrip()[:160])" >> gpurun_out/ablate.log
  done
fi
