cd "$(dirname "$0")/.."
nvidia-smi -L
timeout 1200 python -X faulthandler -m pytest tests/test_gpu_multi.py tests/test_gpu_ref_scripts.py -m gpu -q -x > gpurun_out/r2f_pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/r2f_pytest_multi.log
tail -30 gpurun_out/r2f_pytest_multi.log | cut -c1-300
for halo in peer nccl; do
ECOLAB_B200_HALO=$halo timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2f_bench_n2_$halo.json 2> gpurun_out/r2f_bench_n2_$halo.err; echo "bench2 $halo rc=$?"
tail -3 gpurun_out/r2f_bench_n2_$halo.err | cut -c1-300
cut -c1-2500 gpurun_out/r2f_bench_n2_$halo.json
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2f_bench_ref_n2.json 2> gpurun_out/r2f_bench_ref_n2.err; echo "ref2 rc=$?"
cut -c1-600 gpurun_out/r2f_bench_ref_n2.json
