set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2m_pytest.log
python bench.py --steps 5 --no-cpu > gpurun_out/r2m_bench.json 2> gpurun_out/r2m_bench.err
LZ4E_B200_GIN_MIN=1024 python bench.py --steps 5 --no-e2e --no-cpu --no-replay --no-extra > gpurun_out/r2m_ginmin1024.json 2>&1
tail -n 5 gpurun_out/r2m_pytest.log
