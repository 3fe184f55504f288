cd "$(dirname "$0")/.."
timeout 900 python -X faulthandler -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "generate" > gpurun_out/r2o_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2o_pytest.log
tail -12 gpurun_out/r2o_pytest.log | cut -c1-250
P='import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        j=json.loads(l); k=j["roofline"]["kernels_ms_per_step"]
        print({a:round(b,3) for a,b in k.items()}, "ms/step %.3f"%j["ms_per_step"])
    else: print(l.rstrip()[:300])'
: > gpurun_out/r2o_pair.log
run() { echo "== $*" >> gpurun_out/r2o_pair.log; timeout 300 python bench.py --nx 1024 --ny 1024 --steps 5 --warmup 3 --no-cpu --no-e2e --no-extra --no-verify "$@" 2>&1 | python -c "$P" >> gpurun_out/r2o_pair.log; }
for nnz in 8 16 51; do
run --nnz $nnz --gen-pair -1
run --nnz $nnz
done
run --nnz 3 --gen-nobsmem 1 --gen-pair -1
run --nnz 3 --gen-nobsmem 1
cat gpurun_out/r2o_pair.log
