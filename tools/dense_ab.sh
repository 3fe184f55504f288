# A/B of the dense epilogue's TMEM release point (default: right after the drain; tools/variants/lib_late.so: after the update,
# as in round 1) + the int8 tensor ceiling (tools/ubench_i8peak).  Run under gpurun; output gpurun_out/dn_ab.log, i8peak.txt
mkdir -p gpurun_out; : > gpurun_out/dn_ab.log
for rep in 1 2; do for lib in default late; do for dmax in 199 1000; do
  if [ $lib = default ]; then unset ECOLAB_B200_LIB; else export ECOLAB_B200_LIB=$PWD/tools/variants/lib_late.so; fi
  echo "== $lib dmax=$dmax rep=$rep" >> gpurun_out/dn_ab.log
  timeout 120 python bench.py --dense --dense-prof --no-cpu --steps 40 --warmup 5 --dense-dmax $dmax 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); r=j['roofline']; print(r['kernels_ms_per_step'], r.get('cta0_cycles_last_launch'), j['clocks'])
    else: print(l.rstrip()[:200])" >> gpurun_out/dn_ab.log
done; done; done
unset ECOLAB_B200_LIB
timeout 60 tools/ubench_i8peak > gpurun_out/i8peak.txt 2>&1
