cd "$(dirname "$0")/.."
nvidia-smi -L | wc -l
timeout 1200 python -X faulthandler -m pytest tests/test_gpu_multi.py tests/test_gpu_ref_scripts.py -m gpu -q -x -k "partitioned or two_ranks" > gpurun_out/r2k_pytest_multi8.log 2>&1; echo "rc=$?" >> gpurun_out/r2k_pytest_multi8.log
tail -6 gpurun_out/r2k_pytest_multi8.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29527 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r2k_bench_n8.json 2> gpurun_out/r2k_bench_n8.err; echo "bench8 rc=$?"
tail -3 gpurun_out/r2k_bench_n8.err | cut -c1-300
python - <<'PY'
import json
j=json.load(open('gpurun_out/r2k_bench_n8.json'))
print(j['ms_per_step'], j['value'], j['roofline']['kernels_ms_per_step'], j['gpu_launches'])
print(j['e2e'])
print(j['verify'])
PY
timeout 300 python tools/pcie_probe.py 2>&1 | tail -5
