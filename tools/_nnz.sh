python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "generate or pre_rounding" > gpurun_out/pytest_par.log 2>&1; echo rc=$? >> gpurun_out/pytest_par.log
B="python bench.py --steps 10 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e"
: > gpurun_out/nnz.log
for z in 3 8 16 51; do echo "== nnz $z" >> gpurun_out/nnz.log; $B --nnz $z 2>&1 | grep -o "kernels_ms_per_step[^}]*}" >> gpurun_out/nnz.log; done
