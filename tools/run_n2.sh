run() { p=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $p "$@"; }
run 29501 tests/dist_parity.py --p2p --eval-steps 2 2>&1 | tail -3
run 29502 tests/dist_parity.py --p2p --chunks 2 2>&1 | tail -3
run 29503 tests/dist_parity.py --p2p --chunks 2 --bf16 2>&1 | tail -3
run 29504 bench.py --gpus 2 --steps 20 --warmup 5 --chunks 1 --timeline gpurun_out/tl_n2c1 > gpurun_out/b_n2c1.log 2>&1; tail -c 600 gpurun_out/b_n2c1.log
run 29505 bench.py --gpus 2 --steps 20 --warmup 5 --chunks 2 --timeline gpurun_out/tl_n2c2 > gpurun_out/b_n2c2.log 2>&1; tail -c 600 gpurun_out/b_n2c2.log
