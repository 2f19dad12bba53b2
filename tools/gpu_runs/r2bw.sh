set -x
timeout 600 python -m pytest tests/test_gpu_replay.py -m gpu -x -q > gpurun_out/r2bw_tests.log 2>&1
tail -n 3 gpurun_out/r2bw_tests.log
timeout 900 python bench.py > gpurun_out/r2bw_bench.json 2> gpurun_out/r2bw_bench.err
tail -n 3 gpurun_out/r2bw_bench.err
