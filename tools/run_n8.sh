N=${1:-8}
python -m cloudtik_b200.launcher --nproc-per-node $N --master-port 29521 bench.py --gpus $N --steps 20 --warmup 5 --timeline gpurun_out/tl_n${N} > gpurun_out/b_n${N}.log 2>&1; tail -c 300 gpurun_out/b_n${N}.log
python -m cloudtik_b200.launcher --nproc-per-node $N --master-port 29522 bench.py --gpus $N --steps 20 --warmup 5 --global-batch 65536 --no-c5 > gpurun_out/b_n${N}_strong.log 2>&1; tail -c 200 gpurun_out/b_n${N}_strong.log
