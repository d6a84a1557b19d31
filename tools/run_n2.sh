N=2
L="python -m cloudtik_b200.launcher --nproc-per-node $N"
$L --master-port 29521 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/final_n${N}_weak.log 2>&1; tail -c 100 gpurun_out/final_n${N}_weak.log; echo
$L --master-port 29522 bench.py --gpus $N --steps 20 --warmup 5 --global-batch 65536 --no-c5 --no-parity > gpurun_out/final_n${N}_strong.log 2>&1; tail -c 100 gpurun_out/final_n${N}_strong.log; echo
