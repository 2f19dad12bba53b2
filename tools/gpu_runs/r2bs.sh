set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2bs_n2.json 2> gpurun_out/r2bs_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r2bs_n2_ref.json 2> gpurun_out/r2bs_n2_ref.err
tail -n 3 gpurun_out/r2bs_n2.err gpurun_out/r2bs_n2_ref.err
