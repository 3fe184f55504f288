# A/B of the dense generate kernel on ONE box, alternated: the in-tree library against tools/variants/lib_*.so (built with
# tools/variants.sh, or by hand from an older commit).  Run under gpurun; output gpurun_out/dn_ab.log
mkdir -p gpurun_out; : > gpurun_out/dn_ab.log
for rep in 1 2 3; do for lib in default tools/variants/lib_*.so; do
  if [ $lib = default ]; then unset ECOLAB_B200_LIB; else export ECOLAB_B200_LIB=$PWD/$lib; fi
  echo "== $lib rep=$rep" >> gpurun_out/dn_ab.log
  timeout 120 python bench.py --dense --no-cpu --steps 40 --warmup 5 --dense-dmax ${DMAX:-199} $DENSE_AB_ARGS 2>&1 | python -c '
import sys, json
for l in sys.stdin:
    if l.startswith("{"):
        j = json.loads(l); r = j["roofline"]
        print(r["kernels_ms_per_step"], r.get("cta0_cycles_last_launch"), r.get("fill_bound"), j["clocks"])
    else:
        print(l.rstrip()[:200])' >> gpurun_out/dn_ab.log
done; done
unset ECOLAB_B200_LIB
cut -c1-300 gpurun_out/dn_ab.log
