# Final round-2 validation on one B200: GPU test suite, then the default bench line.
set -x
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/final_gpu_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/final_gpu_tests.log
timeout 300 python bench.py > gpurun_out/final_n1_default.log 2> gpurun_out/final_n1_default.err; echo "bench rc=$?"
python - <<PY
import json
for l in open("gpurun_out/final_n1_default.log"):
    if l.startswith("{"):
        j=json.loads(l); print(j["value"], j["ms_per_step"], j["e2e"]["value"], j["roofline"]["kernel"], j["roofline"]["kernel_ms"], j["roofline"]["frac"], j["final_loss"]); print(j.get("fp32")); print(j["config5"]["value"], j["config5"]["final_loss"])
PY
