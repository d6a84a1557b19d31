N=8
python -m cloudtik_b200.launcher --nproc-per-node $N --master-port 29521 bench.py --workload c5 --gpus $N --steps 10 --warmup 3 > gpurun_out/final_n${N}_c5.log 2>&1; tail -c 400 gpurun_out/final_n${N}_c5.log
