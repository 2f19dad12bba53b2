set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2cd_n8.json 2> gpurun_out/r2cd_n8.err
tail -n 3 gpurun_out/r2cd_n8.err
