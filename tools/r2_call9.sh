cd "$(dirname "$0")/.."
timeout 1500 python -X faulthandler -m pytest tests -m gpu -q -x --deselect tests/test_gpu_multi.py > gpurun_out/r2i_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2i_pytest.log
tail -8 gpurun_out/r2i_pytest.log | cut -c1-250
timeout 300 python tools/cadence_profile.py --mutation 1e-5 --rounds 5 > gpurun_out/r2i_cadence_1e-5.log 2>&1; cat gpurun_out/r2i_cadence_1e-5.log | cut -c1-600
timeout 900 python -X faulthandler bench.py > gpurun_out/r2i_bench_c4.json 2> gpurun_out/r2i_bench_c4.err; echo "bench rc=$?"
tail -3 gpurun_out/r2i_bench_c4.err | cut -c1-300
python - <<'PY'
import json
j=json.load(open('gpurun_out/r2i_bench_c4.json'))
print(j['ms_per_step'], j['roofline']['kernels_ms_per_step'], j['roofline']['frac'], j['roofline']['step_frac_at_16B'])
print(j['e2e']['ms_per_step'], j['verify'])
print(json.dumps(j['extra']['cadence_c2'])[:900])
print(json.dumps(j['extra']['nnz_sweep_c3'])[:900])
PY
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --cache-control none -k regex:'generate_pipe_kernel|migrate_kernel' --launch-skip 6 --launch-count 4 --csv --log-file gpurun_out/r2i_c4_traffic.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-extra --no-verify > gpurun_out/r2i_c4_traffic.log 2>&1; echo "c4 traffic rc=$?"
grep migrate_kernel gpurun_out/r2i_c4_traffic.csv | cut -d, -f5,13- | head -6
