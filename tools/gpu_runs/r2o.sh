set -x
for v in v2 v1 cur; do
  if [ $v = cur ]; then unset LZ4E_B200_LIB; else export LZ4E_B200_LIB=$PWD/variants/lib_$v.so; fi
  python bench.py --steps 5 --no-e2e --no-cpu --no-replay --no-extra > gpurun_out/r2o_$v.json 2>&1
done
