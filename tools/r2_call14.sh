cd "$(dirname "$0")/.."
timeout 2400 python -X faulthandler -m pytest tests -m gpu -q > gpurun_out/r2m_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2m_pytest.log
tail -8 gpurun_out/r2m_pytest.log | cut -c1-300
timeout 900 python -X faulthandler bench.py > gpurun_out/r2m_bench_c4.json 2> gpurun_out/r2m_bench_c4.err; echo "bench rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29537 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2m_bench_n2.json 2> gpurun_out/r2m_bench_n2.err; echo "bench2 rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/r2m_bench_c4.json','gpurun_out/r2m_bench_n2.json'):
    j=json.load(open(f))
    print(f, j['ms_per_step'], j['value'], j['roofline']['kernels_ms_per_step'], j['roofline']['frac'], j['roofline']['step_frac_at_16B'], j['e2e']['ms_per_step'], j['verify']['band_equal_oracle'], j['verify']['species_totals_sha1'])
    if j.get('extra'): print(json.dumps(j['extra']['cadence_c2'])[:700])
PY
