set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2cb_tests.log 2>&1
tail -n 3 gpurun_out/r2cb_tests.log
timeout 600 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2cb_bench.json 2> gpurun_out/r2cb_bench.err
timeout 300 python tools/class_bench.py 65536 3 > gpurun_out/r2cb_class4k.log 2>&1
cat gpurun_out/r2cb_class4k.log
