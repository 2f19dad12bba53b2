set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2i_pytest.log
for m in 1 2 0; do
LZ4E_B200_XFER=$m LZ4E_B200_TRACE=1 timeout 300 python tools/e2e_probe.py > gpurun_out/r2i_e2e_probe_mode$m.log 2>&1
done
for mb in 8 16 32 128; do
LZ4E_B200_CHUNK_MB=$mb LZ4E_B200_TRACE=1 timeout 300 python tools/e2e_probe.py > gpurun_out/r2i_e2e_probe_chunk$mb.log 2>&1
done
python bench.py --steps 5 --no-cpu --no-replay > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err
tail -15 gpurun_out/r2i_pytest.log
tail -4 gpurun_out/r2i_e2e_probe_mode*.log gpurun_out/r2i_e2e_probe_chunk*.log
