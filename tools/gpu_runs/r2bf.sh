set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2bf_tests.log 2>&1
tail -n 3 gpurun_out/r2bf_tests.log
timeout 600 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2bf_bench.json 2> gpurun_out/r2bf_bench.err
tail -c 600 gpurun_out/r2bf_bench.err
cat gpurun_out/r2bf_bench.json
