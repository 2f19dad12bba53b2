set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2v_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2v_pytest.log
python bench.py --steps 5 --no-cpu --no-e2e > gpurun_out/r2v_bench.json 2> gpurun_out/r2v_bench.err
LZ4E_B200_TRACE=1 tools/replay/lz4e_replay --arena-bios --batches 8 > gpurun_out/r2v_replay_trace.log 2>&1
