set -x
for v in a b c; do
  LZ4E_B200_TMA=1 LZ4E_B200_LIB=$PWD/variants/lib_$v.so python bench.py --steps 5 --no-e2e --no-cpu --no-replay --no-extra > gpurun_out/r2q_$v.json 2>&1
done
