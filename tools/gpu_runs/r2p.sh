set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2p_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2p_pytest.log
for t in 0 1 0 1; do
LZ4E_B200_TMA=$t python bench.py --steps 5 --no-e2e --no-cpu --no-replay --no-extra >> gpurun_out/r2p_tma$t.json 2>&1
done
