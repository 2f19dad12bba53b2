set -x
for w in 512 768; do
echo "== g26 win $w" >> gpurun_out/r2bv.log
LZ4E_B200_GIN_WIN=$w LZ4E_B200_LIB=$PWD/variants/g26/liblz4e_b200.so timeout 300 python tools/prof_run.py 4096 32768 3 2>&1 | grep -i "64" >> gpurun_out/r2bv.log
done
echo "== g26 default window" >> gpurun_out/r2bv.log
LZ4E_B200_LIB=$PWD/variants/g26/liblz4e_b200.so timeout 300 python tools/prof_run.py 4096 32768 3 2>&1 | grep -i "64" >> gpurun_out/r2bv.log
echo "== base" >> gpurun_out/r2bv.log
timeout 300 python tools/prof_run.py 4096 32768 3 2>&1 | grep -i "64" >> gpurun_out/r2bv.log
cat gpurun_out/r2bv.log
