set -x
LZ4E_B200_SPARSE_SRC=kernel LZ4E_B200_TRACE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 8 --steps 3 --warmup 3 --no-extra --no-cpu > gpurun_out/r2ab_n8_kernel.json 2> gpurun_out/r2ab_n8_kernel.err
LZ4E_B200_XFER=2 LZ4E_B200_TRACE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 8 --steps 3 --warmup 3 --no-extra --no-cpu > gpurun_out/r2ab_n8_xfer2.json 2> gpurun_out/r2ab_n8_xfer2.err
