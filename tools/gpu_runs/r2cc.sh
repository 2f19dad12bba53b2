set -x
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2cc_plain.json 2> gpurun_out/r2cc_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2cc_launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2cc_ncu_launch.log 2>&1
python tools/prof_run.py 262144 32768 1 > gpurun_out/r2cc_prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'lz4e_(compress|decompress)' -c 8 -f -o gpurun_out/r2cc_full python tools/prof_run.py 262144 32768 1 > gpurun_out/r2cc_ncu_full.log 2>&1
ls -la gpurun_out/
