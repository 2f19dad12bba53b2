set -x
export LZ4E_B200_LIB=$PWD/variants/w10/liblz4e_b200.so
CLASS_SIZE=65536 timeout 600 python tools/class_bench.py 16384 3 > gpurun_out/r2bo_class64k.log 2>&1
cat gpurun_out/r2bo_class64k.log
timeout 300 python tools/prof_run.py 4096 32768 3 2>&1 | grep -i "64" > gpurun_out/r2bo_prof.log
cat gpurun_out/r2bo_prof.log
