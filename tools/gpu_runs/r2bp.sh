set -x
for v in l10 l12; do
LZ4E_B200_LIB=$PWD/variants/$v/liblz4e_b200.so timeout 600 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2bp_$v.json 2> gpurun_out/r2bp_$v.err
done
timeout 600 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2bp_base.json 2> gpurun_out/r2bp_base.err
