set -x
LZ4E_B200_TRACE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2aa_n8.json 2> gpurun_out/r2aa_n8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r2aa_n4.json 2> gpurun_out/r2aa_n4.err
nvidia-smi topo -m > gpurun_out/r2aa_topo.txt 2>&1
lscpu | head -25 > gpurun_out/r2aa_lscpu.txt 2>&1
