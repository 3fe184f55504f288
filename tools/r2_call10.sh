cd "$(dirname "$0")/.."
nvidia-smi topo -m 2>&1 | head -12
timeout 300 python tools/pcie_probe.py 2>&1 | tail -6
timeout 1500 python -X faulthandler -m pytest tests/test_gpu_multi.py tests/test_gpu_ref_scripts.py tests/test_gpu_parity.py -m gpu -q -x -k "partitioned or two_ranks or migrate" > gpurun_out/r2j_pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/r2j_pytest_multi.log
tail -12 gpurun_out/r2j_pytest_multi.log | cut -c1-300
