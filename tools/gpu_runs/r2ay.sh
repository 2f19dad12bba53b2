for v in b2 b8 cur; do
  if [ $v = cur ]; then unset LZ4E_B200_LIB; else export LZ4E_B200_LIB=$PWD/variants/lib_$v.so; fi
  echo "== $v" >> gpurun_out/r2ay.log
  DBENCH_ONLY=0,1,-1,1 python tools/dbench.py 262144 0 5 2>&1 | grep variant >> gpurun_out/r2ay.log
done
