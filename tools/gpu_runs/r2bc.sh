set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2bc_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2bc_pytest.log
tools/replay/lz4e_replay --arena-bios --batches 24 --oracle oracle/_ref/liblz4e_ref.so > gpurun_out/r2bc_one.log 2>&1
tools/replay/lz4e_replay --arena-bios --batches 24 --devices 2 --oracle oracle/_ref/liblz4e_ref.so > gpurun_out/r2bc_two.log 2>&1
