set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2cd_n2.json 2> gpurun_out/r2cd_n2.err
tail -n 3 gpurun_out/r2cd_n2.err
