# round-end check on the GPU box: full GPU test suite, smoke(), default bench, reference arm (run under gpurun)
python -m pytest tests -m gpu -x -q > gpurun_out/val_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/val_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/val_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/val_smoke.log
( time python bench.py ) > gpurun_out/val_bench.json 2> gpurun_out/val_bench.err
( time python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/val_ref.json 2> gpurun_out/val_ref.err
