cd "$(dirname "$0")/.."
timeout 300 python tools/cadence_profile.py --mutation 1e-5 --rounds 5 > gpurun_out/r2h_cadence_1e-5.log 2>&1; cat gpurun_out/r2h_cadence_1e-5.log | cut -c1-600
timeout 300 python tools/cadence_profile.py --mutation 1e-3 --rounds 3 > gpurun_out/r2h_cadence_1e-3.log 2>&1; cat gpurun_out/r2h_cadence_1e-3.log | cut -c1-600
bash tools/ncu_c3.sh r2e
B="python bench.py --steps 3 --warmup 3 --nx 1024 --ny 1024 --no-cpu --no-e2e --no-extra --no-verify --nnz 51"
ncu --set full --clock-control none --import-source on -k regex:generate_pipe_kernel --launch-skip 3 --launch-count 1 -f -o gpurun_out/r2e_generate_nnz51 $B > gpurun_out/r2e_generate_nnz51_ncu.log 2>&1
# C4: DRAM bytes of the two hot kernels in ONE pass (no replay: the 64 GiB buffers are not saved/restored)
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --cache-control none -k regex:'generate_pipe_kernel|migrate_kernel' --launch-skip 6 --launch-count 4 --csv --log-file gpurun_out/r2e_c4_traffic.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-extra --no-verify > gpurun_out/r2e_c4_traffic.log 2>&1; echo "c4 traffic rc=$?"
tail -12 gpurun_out/r2e_c4_traffic.csv | cut -c1-400
ls -la gpurun_out/*.ncu-rep
