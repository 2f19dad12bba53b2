set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2au_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2au_pytest.log
python bench.py --steps 5 --no-e2e --no-cpu --no-replay --no-extra > gpurun_out/r2au_bench.json 2>&1
python tools/class_bench.py 131072 3 2>&1 | grep class > gpurun_out/r2au.log
