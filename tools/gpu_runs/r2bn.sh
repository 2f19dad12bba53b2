set -x
timeout 600 python -m pytest tests/test_gpu_variants.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2bn_tests.log 2>&1
tail -n 2 gpurun_out/r2bn_tests.log
CLASS_SIZE=65536 timeout 600 python tools/class_bench.py 16384 3 > gpurun_out/r2bn_class64k.log 2>&1
cat gpurun_out/r2bn_class64k.log
timeout 300 python tools/prof_run.py 4096 32768 3 2>&1 | grep -i "64" > gpurun_out/r2bn_prof.log
cat gpurun_out/r2bn_prof.log
