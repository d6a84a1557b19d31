run() { p=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $p "$@"; }
run 29504 bench.py --gpus 2 --steps 20 --warmup 5 --no-c5 --timeline gpurun_out/tl_n2g > gpurun_out/b_n2g.log 2>&1; tail -c 400 gpurun_out/b_n2g.log
run 29505 bench.py --gpus 2 --steps 20 --warmup 5 --no-c5 --nccl-tail > gpurun_out/b_n2h.log 2>&1; tail -c 200 gpurun_out/b_n2h.log
