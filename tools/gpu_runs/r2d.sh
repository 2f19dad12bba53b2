set -x
python tools/dbench.py 65536 4096 5 > gpurun_out/r2d_mid.jsonl 2>&1
python tools/dbench.py 8192 512 5 > gpurun_out/r2d_small.jsonl 2>&1
