set -x
python tools/prof_run.py 262144 32768 1 > gpurun_out/r2am_prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'lz4e_(compress|decompress)' -c 8 -f -o gpurun_out/r2am_full python tools/prof_run.py 262144 32768 1 > gpurun_out/r2am_ncu_full.log 2>&1
