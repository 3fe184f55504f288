cd "$(dirname "$0")/.."
P='import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        j=json.loads(l); k=j["roofline"]["kernels_ms_per_step"]
        print({a:round(b,3) for a,b in k.items()}, "ms/step %.3f"%j["ms_per_step"])
    else: print(l.rstrip()[:300])'
: > gpurun_out/r2l_migrows.log
run() { echo "== $*" >> gpurun_out/r2l_migrows.log; timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-extra --no-verify "$@" 2>&1 | python -c "$P" >> gpurun_out/r2l_migrows.log; }
for r in 32 64 128 256 512 1024 4096; do
run --mig-rows $r
done
run
run
cat gpurun_out/r2l_migrows.log
B="python bench.py --steps 3 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e --no-extra --no-verify"
ncu --set full --clock-control none --import-source on -k regex:migrate_kernel --launch-skip 3 --launch-count 1 -f -o gpurun_out/r2l_migrate $B > gpurun_out/r2l_migrate_ncu.log 2>&1
ls -la gpurun_out/r2l_migrate.ncu-rep
