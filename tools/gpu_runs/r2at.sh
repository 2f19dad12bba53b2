set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2at_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2at_pytest.log
DBENCH_ONLY=3,32,0,1 python tools/dbench.py 262144 32768 3 2>&1 | grep compress_ms > gpurun_out/r2at.log
python tools/class_bench.py 131072 3 2>&1 | grep class >> gpurun_out/r2at.log
