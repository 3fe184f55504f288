# parity + C3 kernel timings; output gpurun_out/quick.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
B="python bench.py --steps 10 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e"
P='import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        j=json.loads(l); print(j["roofline"]["kernels_ms_per_step"], "value %.3e"%j["value"])
    else: print(l.rstrip()[:200])'
: > gpurun_out/quick.log
run() { echo "== $*" >> gpurun_out/quick.log; timeout 160 $B "$@" 2>&1 | python -c "$P" >> gpurun_out/quick.log; }
run
run --gen-warps 20
run --gen-warps 12
run --nnz 16
run --nnz 51
python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
