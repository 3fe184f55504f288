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
run --gen-dual 1
run --gen-dual 1 --gen-warps 12
run --gen-dual 1 --nnz 16
run --gen-dual 1 --nnz 51
run --gen-dual 1 --gen-warps 12 --nnz 51
