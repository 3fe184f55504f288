cd "$(dirname "$0")/.."
timeout 900 python -X faulthandler -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -m gpu -q -x -k "generate or random_op" > gpurun_out/r2r_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2r_pytest.log
tail -5 gpurun_out/r2r_pytest.log | cut -c1-250
P='import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        j=json.loads(l); k=j["roofline"]["kernels_ms_per_step"]
        print({a:round(b,3) for a,b in k.items()}, "ms/step %.3f"%j["ms_per_step"])
    else: print(l.rstrip()[:300])'
: > gpurun_out/r2r_alu.log
run() { echo "== $*" >> gpurun_out/r2r_alu.log; timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --no-extra --no-verify "$@" 2>&1 | python -c "$P" >> gpurun_out/r2r_alu.log; }
for nnz in 3 8 16 51; do
run --nx 1024 --ny 1024 --nnz $nnz
done
run
cat gpurun_out/r2r_alu.log
