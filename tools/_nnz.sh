B="python bench.py --steps 10 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e"
: > gpurun_out/nnz.log
for z in 8 16 51; do for d in 0 1; do echo "== nnz $z dual $d" >> gpurun_out/nnz.log; $B --nnz $z --gen-dual $d 2>&1 | grep -o "kernels_ms_per_step[^}]*}" >> gpurun_out/nnz.log; done; done
