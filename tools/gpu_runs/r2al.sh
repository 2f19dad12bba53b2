set -x
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2al_plain.json 2> gpurun_out/r2al_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2al_launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2al_ncu_launch.log 2>&1
python tools/prof_run.py 262144 32768 0 > gpurun_out/r2al_prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'lz4e_(compress|decompress)' -c 4 -f -o gpurun_out/r2al_full python tools/prof_run.py 262144 32768 0 > gpurun_out/r2al_ncu_full.log 2>&1
