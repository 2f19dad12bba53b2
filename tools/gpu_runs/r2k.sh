set -x
for m in kernel hull; do
LZ4E_B200_SPARSE_SRC=$m LZ4E_B200_TRACE=2 timeout 300 python tools/e2e_probe.py > gpurun_out/r2k_e2e_probe_$m.log 2>&1
done
LZ4E_B200_XFER=0 LZ4E_B200_TRACE=2 timeout 300 python tools/e2e_probe.py > gpurun_out/r2k_e2e_probe_xfer0.log 2>&1
