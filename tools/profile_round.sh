# ncu --set full of the two hot kernels at C3, the C4 launch list and the aux-kernel table (run under gpurun); tag = r1c
bash tools/ncu_c3.sh r1c
python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r1c_c4_plain.json 2> gpurun_out/r1c_c4_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1c_launches_c4.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r1c_c4_ncu.log 2>&1
python tools/bench_aux.py > gpurun_out/r1c_aux.json 2> gpurun_out/r1c_aux.err
