B="python bench.py --steps 10 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e"
: > gpurun_out/span.log
for t in 1 2 4 8; do echo "== span $t" >> gpurun_out/span.log; $B --mig-variant $t 2>&1 | grep -o "kernels_ms_per_step[^}]*}" >> gpurun_out/span.log; done
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "span" > gpurun_out/pytest_span.log 2>&1
