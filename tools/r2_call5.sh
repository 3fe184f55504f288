cd "$(dirname "$0")/.."
timeout 1500 python -X faulthandler -m pytest tests -m gpu -q --deselect tests/test_gpu_multi.py > gpurun_out/r2e_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2e_pytest.log
tail -40 gpurun_out/r2e_pytest.log | cut -c1-250
timeout 900 python -X faulthandler bench.py > gpurun_out/r2e_bench_c4.json 2> gpurun_out/r2e_bench_c4.err; echo "bench rc=$?"
tail -5 gpurun_out/r2e_bench_c4.err | cut -c1-300
cat gpurun_out/r2e_bench_c4.json | cut -c1-6000
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2e_bench_ref.json 2> gpurun_out/r2e_bench_ref.err; echo "ref rc=$?"
cat gpurun_out/r2e_bench_ref.json | cut -c1-1500
nvidia-smi --query-gpu=name,memory.total --format=csv
nproc; free -g | head -2
