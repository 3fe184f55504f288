cd "$(dirname "$0")/.."
timeout 1500 python -X faulthandler -m pytest tests -m gpu -q --deselect tests/test_gpu_multi.py > gpurun_out/r2c_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2c_pytest.log
tail -30 gpurun_out/r2c_pytest.log
P='import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        j=json.loads(l); k=j["roofline"]["kernels_ms_per_step"]
        print({a:round(b,3) for a,b in k.items()}, "ms/step %.3f"%j["ms_per_step"])
    else: print(l.rstrip()[:300])'
: > gpurun_out/r2c_walker.log
run() { echo "== $*" >> gpurun_out/r2c_walker.log; timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-extra --no-verify "$@" 2>&1 | python -c "$P" >> gpurun_out/r2c_walker.log; }
for i in 1 2 3; do
run --mig-variant 1
run --nx 4112 --ny 4080 --mig-variant 1
run --nx 4128 --ny 4064 --mig-variant 1
done
cat gpurun_out/r2c_walker.log
# isolate the crash of the full bench line: one leg at a time, Python traceback on SIGSEGV
timeout 600 python -X faulthandler bench.py --nx 1024 --ny 1024 --steps 3 --no-cpu --no-e2e --no-extra > gpurun_out/r2c_b_verify.json 2> gpurun_out/r2c_b_verify.err; echo "verify rc=$?"
timeout 600 python -X faulthandler bench.py --nx 1024 --ny 1024 --steps 3 --no-cpu --no-verify --no-extra --e2e-n 1024 > gpurun_out/r2c_b_e2e.json 2> gpurun_out/r2c_b_e2e.err; echo "e2e rc=$?"
timeout 600 python -X faulthandler bench.py --nx 1024 --ny 1024 --steps 3 --no-cpu --no-verify --no-e2e --no-dense > gpurun_out/r2c_b_extra.json 2> gpurun_out/r2c_b_extra.err; echo "extra rc=$?"
timeout 600 python -X faulthandler bench.py --nx 1024 --ny 1024 --steps 3 --no-verify --no-e2e --no-extra > gpurun_out/r2c_b_cpu.json 2> gpurun_out/r2c_b_cpu.err; echo "cpu rc=$?"
for f in verify e2e extra cpu; do echo "--- $f"; tail -25 gpurun_out/r2c_b_$f.err | cut -c1-250; done
