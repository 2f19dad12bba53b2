set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2ag_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2ag_pytest.log
LZ4E_B200_TRACE=1 timeout 300 python tools/e2e_probe.py > gpurun_out/r2ag_e2e_probe.log 2>&1
python bench.py --steps 5 --no-cpu --no-extra > gpurun_out/r2ag_bench.json 2> gpurun_out/r2ag_bench.err
