set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2af_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2af_pytest.log
LZ4E_B200_TRACE=1 python bench.py --steps 5 --no-cpu --no-extra > gpurun_out/r2af_bench.json 2> gpurun_out/r2af_bench.err
LZ4E_B200_SPARSE_SRC=kernel python bench.py --steps 5 --no-cpu --no-extra > gpurun_out/r2af_bench_kernel.json 2> gpurun_out/r2af_bench_kernel.err
