set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2bb_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2bb_pytest.log
DBENCH_ONLY=3,32,0,1 python tools/dbench.py 262144 32768 5 2>&1 | grep variant > gpurun_out/r2bb.log
DBENCH_ONLY=3,32,0,1 python tools/dbench.py 0 512 5 2>&1 | grep variant >> gpurun_out/r2bb.log
