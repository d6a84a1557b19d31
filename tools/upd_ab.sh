# window_update_kernel: one chunk per block (0) vs one wave of chunk-claiming blocks with the join deferred (auto),
# flush depth 16 / 4
set -x
B200_UPD_PERSIST=0 timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/upd_tests_0.log 2>&1; echo "rc0=$?"; tail -2 gpurun_out/upd_tests_0.log
B200_UPD_PERSIST=1 B200_UPD_GRID=3 B200_UPD_KFLUSH=2 timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/upd_tests_g3.log 2>&1; echo "rcg3=$?"; tail -2 gpurun_out/upd_tests_g3.log
run() {
B200_UPD_PERSIST=$1 B200_UPD_KFLUSH=$2 timeout 200 python bench.py --steps 10 --warmup 3 --no-c5 --no-microbench --no-cpu-baseline --no-fp32 > gpurun_out/upd_bench_$1_$2.log 2>&1; echo "bench $1 $2 rc=$?"
python - <<PY
import json
for l in open("gpurun_out/upd_bench_$1_$2.log"):
    if l.startswith("{"):
        j=json.loads(l); print("$1 $2", j["value"], j["ms_per_step"], j["roofline"]["kernel"], j["roofline"]["kernel_ms"], j["roofline"]["frac"], j["final_loss"])
PY
}
run 0 16
run auto 16
run auto 4
