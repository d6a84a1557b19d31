N=${1:-8}
L="python -m cloudtik_b200.launcher --nproc-per-node $N"
$L --master-port 29521 bench.py --gpus $N --steps 20 --warmup 5 --no-c5 --no-parity --timeline gpurun_out/tl_n${N}d > gpurun_out/b_n${N}_e.log 2>&1; tail -c 100 gpurun_out/b_n${N}_e.log; echo
B200_FWD_CTAS=4 $L --master-port 29522 bench.py --gpus $N --steps 20 --warmup 5 --no-c5 --no-parity > gpurun_out/b_n${N}_f.log 2>&1; tail -c 100 gpurun_out/b_n${N}_f.log; echo
B200_FWD_CTAS=2 $L --master-port 29523 bench.py --gpus $N --steps 20 --warmup 5 --no-c5 --no-parity > gpurun_out/b_n${N}_g.log 2>&1; tail -c 100 gpurun_out/b_n${N}_g.log; echo
