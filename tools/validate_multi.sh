# round-end check on a multi-GPU box (run under gpurun --gpus N): partitioned parity tests, the reference's scripts on two
# ranks, the default bench on all GPUs and the reference arm under torchrun
cd "$(dirname "$0")/.."
N=$(nvidia-smi -L | wc -l)
timeout 1500 python -X faulthandler -m pytest tests/test_gpu_multi.py tests/test_gpu_ref_scripts.py -m gpu -q -k "partitioned or two_ranks" > gpurun_out/valm_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/valm_pytest.log
tail -6 gpurun_out/valm_pytest.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29547 bench.py --gpus $N > gpurun_out/valm_bench_n$N.json 2> gpurun_out/valm_bench_n$N.err; echo "bench$N rc=$?"
tail -2 gpurun_out/valm_bench_n$N.err | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29549 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/valm_ref_n$N.json 2> gpurun_out/valm_ref_n$N.err; echo "ref$N rc=$?"
python - <<PY
import json
j=json.load(open('gpurun_out/valm_bench_n$N.json'))
print(j['ms_per_step'], j['value'], j['roofline']['kernels_ms_per_step'], j['gpu_launches'], j['roofline']['step_frac_at_16B'])
print(j['e2e'])
print(j['verify'])
r=json.load(open('gpurun_out/valm_ref_n$N.json'))
print(r['value'], r['cpu_baseline']['cores'])
PY
