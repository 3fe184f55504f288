cd "$(dirname "$0")/.."
free -g | head -2
time timeout 900 python -X faulthandler bench.py --no-cpu --no-extra --no-verify --steps 5 > gpurun_out/r2p_bench_e2e.json 2> gpurun_out/r2p_bench_e2e.err; echo "bench rc=$?"
true
python - <<'PY'
import json
j=json.load(open('gpurun_out/r2p_bench_e2e.json'))
print(j['ms_per_step'], j['e2e'])
PY
