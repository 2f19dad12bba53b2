set -x
for w in 0 14 32 48; do
echo "== wave_serial $w" >> gpurun_out/r2ae.log
LZ4E_B200_WAVE_SERIAL=$w CLASS_SIZE=65536 python tools/class_bench.py 8192 3 >> gpurun_out/r2ae.log 2>&1
LZ4E_B200_WAVE_SERIAL=$w DBENCH_ONLY=3,32,0,1 python tools/dbench.py 0 32768 5 2>&1 | grep variant >> gpurun_out/r2ae.log
done
