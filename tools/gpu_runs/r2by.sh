set -x
export LZ4E_B200_LIB=$PWD/variants/pin/liblz4e_b200.so
timeout 600 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2by_bench.json 2> gpurun_out/r2by_bench.err
timeout 300 python tools/class_bench.py 65536 3 > gpurun_out/r2by_class4k.log 2>&1
cat gpurun_out/r2by_class4k.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_variants.py -m gpu -x -q > gpurun_out/r2by_tests.log 2>&1
tail -n 3 gpurun_out/r2by_tests.log
