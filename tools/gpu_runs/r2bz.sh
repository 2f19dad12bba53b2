set -x
for r in 1 2; do for v in pin pin2; do
LZ4E_B200_LIB=$PWD/variants/$v/liblz4e_b200.so timeout 600 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2bz_${v}_$r.json 2> gpurun_out/r2bz_${v}_$r.err
done; done
