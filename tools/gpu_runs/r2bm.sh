set -x
timeout 300 python tools/class_bench.py 65536 3 > gpurun_out/r2bm_class4k.log 2>&1
cat gpurun_out/r2bm_class4k.log
