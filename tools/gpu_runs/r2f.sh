set -x
export DBENCH_ONLY=3,32,0,1
ncu --set full --clock-control none --import-source on -k regex:lz4e_decompress_wave -c 6 -f -o gpurun_out/r2f_wave python tools/dbench.py 262144 32768 1 > gpurun_out/r2f_ncu.log 2>&1
tail -3 gpurun_out/r2f_ncu.log
