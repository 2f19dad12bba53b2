set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2j_pytest.log
for t in 4 8 2; do
LZ4E_B200_HOST_THREADS=$t LZ4E_B200_TRACE=1 timeout 300 python tools/e2e_probe.py > gpurun_out/r2j_e2e_probe_t$t.log 2>&1
done
for mb in 32 64; do
LZ4E_B200_CHUNK_MB=$mb LZ4E_B200_TRACE=1 timeout 300 python tools/e2e_probe.py > gpurun_out/r2j_e2e_probe_chunk$mb.log 2>&1
done
python bench.py --steps 5 --no-cpu --no-replay > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err
tail -n 15 gpurun_out/r2j_pytest.log
