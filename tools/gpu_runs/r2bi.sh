set -x
CLASS_SIZE=65536 timeout 600 python tools/class_bench.py 16384 3 > gpurun_out/r2bi_class64k.log 2>&1
cat gpurun_out/r2bi_class64k.log
