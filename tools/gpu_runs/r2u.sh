set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2u_pytest.log
tools/replay/lz4e_replay --arena-bios --batches 24 > gpurun_out/r2u_replay.log 2>&1
LZ4E_B200_TRACE=1 tools/replay/lz4e_replay --arena-bios --batches 8 > gpurun_out/r2u_replay_trace.log 2>&1
