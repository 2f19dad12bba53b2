set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
timeout 600 python tools/dbench.py > gpurun_out/r2b_dbench.jsonl 2> gpurun_out/r2b_dbench.err
tail -3 gpurun_out/r2b_pytest.log
