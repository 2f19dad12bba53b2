set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2as_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2as_pytest.log
DBENCH_ONLY=3,32,0,1 python tools/dbench.py 262144 32768 3 2>&1 | grep compress_ms > gpurun_out/r2as.log
CLASS_SIZE=65536 python tools/class_bench.py 8192 3 2>&1 | grep class >> gpurun_out/r2as.log
tools/replay/lz4e_replay --arena-bios --batches 24 2>&1 | tail -n 1 | cut -c100-240 >> gpurun_out/r2as.log
