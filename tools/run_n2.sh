run() { p=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $p "$@"; }
run 29501 tests/dist_parity.py --p2p --rowwise 2 2>&1 | tail -1
run 29502 tests/dist_parity.py --p2p --rowwise 2 --adagrad --bf16 2>&1 | tail -1
run 29504 tests/dist_parity.py --p2p --boundary 1000 2>&1 | tail -1
run 29503 bench.py --workload c5 --gpus 2 --steps 10 --warmup 3 > gpurun_out/b_n2_c5.log 2>&1; tail -c 900 gpurun_out/b_n2_c5.log
