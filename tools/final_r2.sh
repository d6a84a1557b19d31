# driver entry points, last check of the round: smoke() and the reference arm
set -x
timeout 100 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/final_smoke.log
timeout 100 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/final_ref.log 2>&1; echo "ref rc=$?"; tail -c 700 gpurun_out/final_ref.log
