cd "$(dirname "$0")/.."
timeout 1500 python -X faulthandler -m pytest tests -m gpu -q --deselect tests/test_gpu_multi.py > gpurun_out/r2d_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2d_pytest.log
tail -40 gpurun_out/r2d_pytest.log | cut -c1-250
P='import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        j=json.loads(l); k=j["roofline"]["kernels_ms_per_step"]
        print({a:round(b,3) for a,b in k.items()}, "ms/step %.3f"%j["ms_per_step"])
    else: print(l.rstrip()[:300])'
: > gpurun_out/r2d_walker.log
run() { echo "== $*" >> gpurun_out/r2d_walker.log; timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-extra --no-verify "$@" 2>&1 | python -c "$P" >> gpurun_out/r2d_walker.log; }
for r in 16 32 64 128 256 512 1366 4096; do
run --mig-variant 1 --mig-rows $r
done
for r in 32 64 128; do
run --mig-variant 1 --mig-rows $r
done
run --mig-variant 2 --mig-rows 64
run --mig-variant 2 --mig-rows 4096
cat gpurun_out/r2d_walker.log
timeout 900 python -X faulthandler bench.py > gpurun_out/r2d_bench_c4.json 2> gpurun_out/r2d_bench_c4.err; echo "bench rc=$?"
tail -5 gpurun_out/r2d_bench_c4.err | cut -c1-300
