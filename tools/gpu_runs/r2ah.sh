set -x
CLASS_ONLY=0 ncu --set full --clock-control none --import-source on -k regex:lz4e_compress_kernel -c 1 -s 1 -f -o gpurun_out/r2ah_class0 python tools/class_bench.py 65536 1 > gpurun_out/r2ah_ncu0.log 2>&1
