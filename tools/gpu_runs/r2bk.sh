set -x
for p in 0 6 8 10 12 14 20 24; do
echo "== GIN_PER_SM=$p" >> gpurun_out/r2bk_gin_per_sm.log
LZ4E_B200_GIN_PER_SM=$p timeout 300 python tools/prof_run.py 4096 32768 3 2>&1 | grep -i "64" >> gpurun_out/r2bk_gin_per_sm.log
done
cat gpurun_out/r2bk_gin_per_sm.log
