set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2ce_tests.log 2>&1
tail -n 3 gpurun_out/r2ce_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2ce_smoke.log 2>&1
tail -n 2 gpurun_out/r2ce_smoke.log
timeout 600 python bench.py --steps 5 --no-cpu --no-replay > gpurun_out/r2ce_bench.json 2> gpurun_out/r2ce_bench.err
