N=4
L="python -m cloudtik_b200.launcher --nproc-per-node $N"
B200_FWD_PERSIST=0 $L --master-port 29521 bench.py --workload c5 --gpus $N --steps 2 --warmup 3 > gpurun_out/c5_v1.log 2>&1
$L --master-port 29522 bench.py --workload c5 --gpus $N --steps 2 --warmup 3 --no-graph > gpurun_out/c5_v2.log 2>&1
$L --master-port 29523 bench.py --workload c5 --gpus $N --steps 2 --warmup 3 --pooling 3 > gpurun_out/c5_v3.log 2>&1
for v in 1 2 3; do python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/c5_v$v.log") if x.startswith("{")][0]; d=json.loads(l); print("v$v", d["final_loss"], d["ms_per_step"])
except Exception as e:
    print("v$v ERR", open("gpurun_out/c5_v$v.log").read()[-600:])
PY
done
