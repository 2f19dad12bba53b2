set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2bq_tests.log 2>&1
tail -n 2 gpurun_out/r2bq_tests.log
for p in 0 4 5 6 7 8; do
LZ4E_B200_LANE_PER_SM=$p timeout 600 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2bq_lane$p.json 2> gpurun_out/r2bq_lane$p.err
done
