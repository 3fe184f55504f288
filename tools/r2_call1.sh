# round 2, first GPU call: parity suite + kernel timings of the new migrate pipeline / FMA generate at C3 and C4
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_multi.py > gpurun_out/r2a_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2a_pytest.log
P='import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        j=json.loads(l); print(j["roofline"]["kernels_ms_per_step"], "value %.3e"%j["value"], "ms/step %.3f"%j["ms_per_step"])
    else: print(l.rstrip()[:300])'
: > gpurun_out/r2a_quick.log
B3="python bench.py --steps 10 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e --no-extra --no-verify"
B4="python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-extra --no-verify"
run() { echo "== $*" >> gpurun_out/r2a_quick.log; timeout 300 "$@" 2>&1 | python -c "$P" >> gpurun_out/r2a_quick.log; }
run $B3
run $B3 --mig-variant 1
run $B3 --mig-stages 4
run $B3 --mig-stages 12
run $B3 --gen-dual -1
run $B3 --nnz 16
run $B3 --nnz 51
run $B4
run $B4
run $B4 --mig-variant 1
run $B4 --mig-stages 12
run $B4 --mig-stages 5
timeout 900 python bench.py > gpurun_out/r2a_bench_c4.json 2> gpurun_out/r2a_bench_c4.err; echo "rc=$?" >> gpurun_out/r2a_bench_c4.err
tail -5 gpurun_out/r2a_pytest.log; cat gpurun_out/r2a_quick.log; tail -c 600 gpurun_out/r2a_bench_c4.err
