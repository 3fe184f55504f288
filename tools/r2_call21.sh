cd "$(dirname "$0")/.."
timeout 600 python -X faulthandler -m pytest tests/test_gpu_dense.py -m gpu -q -x > gpurun_out/r2s_dense_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2s_dense_pytest.log
tail -4 gpurun_out/r2s_dense_pytest.log | cut -c1-250
timeout 300 python bench.py --dense --dense-prof --steps 3 --warmup 2 --no-cpu > gpurun_out/r2s_dense_prof.json 2> gpurun_out/r2s_dense_prof.err; echo rc=$?
python - <<'PY'
import json
j=json.load(open('gpurun_out/r2s_dense_prof.json'))
print(j['ms_per_step'], j['roofline']['kernels_ms_per_step'], j['roofline'].get('cta0_cycles_last_launch'))
PY
timeout 300 python bench.py --dense --dense-prof --dense-dmax 60000 --steps 3 --warmup 2 --no-cpu > gpurun_out/r2s_dense_prof_q2.json 2> gpurun_out/r2s_dense_prof_q2.err; echo rc=$?
python - <<'PY'
import json
j=json.load(open('gpurun_out/r2s_dense_prof_q2.json'))
print(j['ms_per_step'], j['roofline']['kernels_ms_per_step'], j['roofline'].get('cta0_cycles_last_launch'))
PY
