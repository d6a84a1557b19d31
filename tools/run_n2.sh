N=2
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 --no-c5 > gpurun_out/final2_n2_weak.log 2> gpurun_out/final2_n2_weak.err
echo "rc=$?"
python - <<PY
import json
for l in open("gpurun_out/final2_n2_weak.log"):
    if l.startswith("{"):
        j=json.loads(l); print(j["value"], j["ms_per_step"], j["parity"]["ok"], j["parity"]["fp32"], j["roofline"]["kernel"], j["roofline"]["kernel_ms"], j["final_loss"])
PY
tail -3 gpurun_out/final2_n2_weak.err
