for g in 24 20 19 16 12; do
echo "gin_per_sm $g" >> gpurun_out/r2ar.log
LZ4E_B200_GIN_PER_SM=$g DBENCH_ONLY=3,32,0,1 python tools/dbench.py 0 32768 3 2>&1 | grep compress_ms >> gpurun_out/r2ar.log
LZ4E_B200_GIN_PER_SM=$g CLASS_SIZE=65536 CLASS_ONLY=0 python tools/class_bench.py 8192 3 2>&1 | grep class >> gpurun_out/r2ar.log
done
