set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2ca_tests.log 2>&1
tail -n 3 gpurun_out/r2ca_tests.log
timeout 900 python bench.py > gpurun_out/r2ca_bench.json 2> gpurun_out/r2ca_bench.err
timeout 900 python bench.py --impl reference > gpurun_out/r2ca_ref.json 2> gpurun_out/r2ca_ref.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2ca_smoke.log 2>&1
tail -n 2 gpurun_out/r2ca_smoke.log
