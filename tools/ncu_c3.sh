# ncu --set full captures of the two hot kernels at C3 (1024 x 1024 cells x 1024 species); run under gpurun.
# usage: bash tools/ncu_c3.sh <tag> [bench args]     -> gpurun_out/<tag>_{migrate,generate}.ncu-rep
tag=$1; shift
B="python bench.py --steps 3 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e --no-extra --no-verify $*"
$B > gpurun_out/${tag}_plain.log 2>&1 || exit 1
for k in migrate generate_pipe; do
  ncu --set full --clock-control none --import-source on -k regex:${k}_kernel --launch-skip 3 --launch-count 1 -f \
      -o gpurun_out/${tag}_${k} $B > gpurun_out/${tag}_${k}_ncu.log 2>&1
done
