set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2ba_smoke.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2ba_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2ba_pytest.log
python bench.py --steps 5 > gpurun_out/r2ba_bench.json 2> gpurun_out/r2ba_bench.err
