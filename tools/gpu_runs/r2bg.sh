set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2bg_tests.log 2>&1
tail -n 3 gpurun_out/r2bg_tests.log
for i in 1 2; do
timeout 600 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2bg_bench$i.json 2> gpurun_out/r2bg_bench$i.err
done
