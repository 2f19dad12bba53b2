set -x
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2bh_alone.log 2>&1
tail -n 3 gpurun_out/r2bh_alone.log
