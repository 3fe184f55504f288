# knob sweep at C3 (1024 x 1024 cells x 1024 species); output gpurun_out/sweep.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu2.log
B="python bench.py --steps 10 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e"
P='import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        j=json.loads(l); print(j["roofline"]["kernels_ms_per_step"], "value %.3e"%j["value"])
    else: print(l.rstrip()[:200])'
: > gpurun_out/sweep.log
run() { echo "== $*" >> gpurun_out/sweep.log; timeout 120 $B "$@" 2>&1 | python -c "$P" >> gpurun_out/sweep.log; }
run
run --mig-variant 1
