python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu2.log
B="python bench.py --steps 10 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e"
: > gpurun_out/sweep.log
for nw in 8 12 16 20 24; do for st in 2 3; do for nf in 0 1; do
  echo "== nw=$nw stages=$st nofma=$nf" >> gpurun_out/sweep.log
  timeout 120 $B --gen-warps $nw --gen-stages $st --gen-nofma $nf 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print(j['roofline']['kernels_ms_per_step'], 'value %.3e'%j['value'])
    else: print(l.rstrip()[:200])" >> gpurun_out/sweep.log
done; done; done
for nnz in 16 51; do for nf in 0 1; do echo "== nnz=$nnz nofma=$nf" >> gpurun_out/sweep.log; timeout 120 $B --nnz $nnz --gen-nofma $nf 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print(j['roofline']['kernels_ms_per_step'], 'value %.3e'%j['value'])
    else: print(l.rstrip()[:200])" >> gpurun_out/sweep.log; done; done
