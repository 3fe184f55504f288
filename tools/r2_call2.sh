# hypothesis: power-of-two row strides put all rows in flight on the same DRAM banks.  Zero-code test: grid widths whose
# tile count per row is odd / not a power of two.
cd "$(dirname "$0")/.."
P='import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        j=json.loads(l); k=j["roofline"]["kernels_ms_per_step"]; u=j["value"]*j["ms_per_step"]*1e-3
        print({a:round(b,3) for a,b in k.items()}, "ns/Mupd gen %.4f mig %.4f"%(k["generate"]*1e6/u*1e-3*1e3, k["migrate"]*1e6/u*1e-3*1e3), "updates %.3e"%u)
    else: print(l.rstrip()[:300])'
: > gpurun_out/r2b_stride.log
run() { echo "== $*" >> gpurun_out/r2b_stride.log; timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-extra --no-verify "$@" 2>&1 | python -c "$P" >> gpurun_out/r2b_stride.log; }
for v in 2 1; do
run --nx 1024 --ny 1024 --mig-variant $v
run --nx 1040 --ny 1008 --mig-variant $v
run --nx 1056 --ny 992 --mig-variant $v
run --nx 1072 --ny 978 --mig-variant $v
run --nx 1008 --ny 1040 --mig-variant $v
done
run --nx 1040 --ny 1008 --mig-variant 2 --mig-stages 12
run --nx 1040 --ny 1008 --mig-variant 2 --mig-stages 4
run --nx 4112 --ny 4080 --mig-variant 2
run --nx 4112 --ny 4080 --mig-variant 1
run --nx 4112 --ny 4080 --mig-variant 2 --mig-stages 12
cat gpurun_out/r2b_stride.log
