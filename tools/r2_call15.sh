cd "$(dirname "$0")/.."
timeout 900 python -X faulthandler -m pytest tests/test_gpu_ref_scripts.py tests/test_gpu_multi.py -m gpu -q -k "two_ranks or 2-" > gpurun_out/r2n_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2n_pytest.log
tail -6 gpurun_out/r2n_pytest.log | cut -c1-300
timeout 300 python tools/cadence_profile.py --mutation 1e-5 --rounds 4 --passes 2 --prealloc-gb 120 > gpurun_out/r2n_cadence.log 2>&1; cut -c1-420 gpurun_out/r2n_cadence.log
