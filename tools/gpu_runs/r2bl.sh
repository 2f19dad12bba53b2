set -x
for c in 0 2 3; do
CLASS_ONLY=$c CLASS_SIZE=65536 LZ4E_B200_LIB=$PWD/variants/nocopy/liblz4e_b200.so timeout 300 python tools/class_bench.py 16384 3 >> gpurun_out/r2bl_nocopy.log 2>&1
done
cat gpurun_out/r2bl_nocopy.log
