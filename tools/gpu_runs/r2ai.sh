set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2ai_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2ai_pytest.log
for p in 1 0 1 0; do
LZ4E_B200_P12=$p python bench.py --steps 5 --no-e2e --no-cpu --no-replay --no-extra >> gpurun_out/r2ai_p12_$p.json 2>&1
done
LZ4E_B200_P12=1 python tools/class_bench.py 131072 3 > gpurun_out/r2ai_class_p12.log 2>&1
