set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2n_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2n_pytest.log
for t in 0 1 0 1; do
LZ4E_B200_TMA=$t python bench.py --steps 5 --no-e2e --no-cpu --no-replay --no-extra >> gpurun_out/r2n_tma$t.json 2>&1
done
LZ4E_B200_TRACE=1 timeout 300 python tools/e2e_probe.py > gpurun_out/r2n_e2e_probe.log 2>&1
tail -n 5 gpurun_out/r2n_pytest.log
