set -x
LZ4E_B200_TRACE=1 tools/replay/lz4e_replay --arena-bios --batches 12 > gpurun_out/r2s_replay.log 2>&1
