set -x
export DBENCH_ONLY=2,8,0
python tools/dbench.py 262144 32768 1 > gpurun_out/r2c_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:lz4e_decompress_group -c 2 -s 3 -f -o gpurun_out/r2c_dgroup python tools/dbench.py 262144 32768 1 > gpurun_out/r2c_ncu.log 2>&1
tail -3 gpurun_out/r2c_plain.log
