set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2h_pytest.log
LZ4E_B200_TRACE=1 timeout 300 python tools/e2e_probe.py > gpurun_out/r2h_e2e_probe.log 2>&1
timeout 300 python tools/dbench.py 0 4096 5 > gpurun_out/r2h_dbench_mid.jsonl 2>&1
timeout 300 python tools/dbench.py 0 512 5 > gpurun_out/r2h_dbench_small.jsonl 2>&1
timeout 300 python tools/dbench.py 0 128 5 > gpurun_out/r2h_dbench_tiny.jsonl 2>&1
tail -5 gpurun_out/r2h_pytest.log
cat gpurun_out/r2h_e2e_probe.log
