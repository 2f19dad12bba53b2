for i in 1 2; do python bench.py --steps 5 --no-e2e --no-cpu --no-replay --no-extra 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['decompress_gbs'], d['kernel_ms'])" >> gpurun_out/r2av.log; done
DBENCH_ONLY=0,1,-1,1 python tools/dbench.py 262144 0 5 2>&1 | grep variant >> gpurun_out/r2av.log
DBENCH_ONLY=0,1,-1,0 python tools/dbench.py 262144 0 5 2>&1 | grep variant >> gpurun_out/r2av.log
