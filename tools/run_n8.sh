N=${1:-8}
run() { p=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $p "$@"; }
run 29504 bench.py --gpus $N --steps 20 --warmup 5 --timeline gpurun_out/tl_n${N} > gpurun_out/b_n${N}.log 2>&1; tail -c 300 gpurun_out/b_n${N}.log
