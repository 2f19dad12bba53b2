set -x
for v in w9 w10p; do
echo "== $v" >> gpurun_out/r2bx.log
LZ4E_B200_LIB=$PWD/variants/$v/liblz4e_b200.so CLASS_SIZE=65536 timeout 600 python tools/class_bench.py 16384 3 >> gpurun_out/r2bx.log 2>&1
LZ4E_B200_LIB=$PWD/variants/$v/liblz4e_b200.so timeout 300 python tools/prof_run.py 4096 32768 3 2>&1 | grep -i "64" >> gpurun_out/r2bx.log
done
cat gpurun_out/r2bx.log
