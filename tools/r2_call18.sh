cd "$(dirname "$0")/.."
timeout 1200 python -X faulthandler -m pytest tests/test_gpu_fuzz.py -m gpu -q > gpurun_out/r2q_fuzz.log 2>&1; echo "rc=$?" >> gpurun_out/r2q_fuzz.log
tail -25 gpurun_out/r2q_fuzz.log | cut -c1-300
timeout 900 python -X faulthandler bench.py > gpurun_out/r2q_bench_c4.json 2> gpurun_out/r2q_bench_c4.err; echo "bench rc=$?"
tail -3 gpurun_out/r2q_bench_c4.err | cut -c1-300
python - <<'PY'
import json
j=json.load(open('gpurun_out/r2q_bench_c4.json'))
print(j['ms_per_step'], j['value'], j['roofline']['kernels_ms_per_step'], j['roofline']['frac'], j['roofline']['step_frac_at_16B'])
print(j['e2e'])
print(json.dumps(j['extra']['cadence_c2'])[:800])
print(j['extra']['dense_c5']['ms_per_step'], j['extra']['dense_c5']['roofline']['frac'])
PY
