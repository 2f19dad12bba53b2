set -x
for g in 2 1 0; do
LZ4E_B200_GIN_SMALL=$g tools/replay/lz4e_replay --arena-bios --batches 24 > gpurun_out/r2t_replay_g$g.log 2>&1
done
LZ4E_B200_GIN_SMALL=1 LZ4E_B200_TRACE=1 tools/replay/lz4e_replay --arena-bios --batches 8 > gpurun_out/r2t_replay_trace.log 2>&1
