set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2ap_n8.json 2> gpurun_out/r2ap_n8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > gpurun_out/r2ap_n8_ref.json 2> gpurun_out/r2ap_n8_ref.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2ap_n4.json 2> gpurun_out/r2ap_n4.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2ap_n2.json 2> gpurun_out/r2ap_n2.err
