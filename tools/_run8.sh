bash tools/variants.sh run
B="python bench.py --steps 10 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e"
for t in 2 4; do echo "== default span $t" >> gpurun_out/variants.log; $B --mig-variant $t 2>&1 | grep -o "kernels_ms_per_step[^}]*}" >> gpurun_out/variants.log; done
