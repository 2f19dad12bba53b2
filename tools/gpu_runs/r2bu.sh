set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2bu_tests.log 2>&1
tail -n 15 gpurun_out/r2bu_tests.log
timeout 600 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2bu_bench.json 2> gpurun_out/r2bu_bench.err
