# Final round-2 validation on one B200: GPU test suite, the default bench line, then (only after the same command exited 0
# without ncu) a launch list and one --set full capture of the two embedding kernels of the eager step.
set -x
timeout 330 python -m pytest tests -m gpu -x -q > gpurun_out/final_gpu_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/final_gpu_tests.log
timeout 300 python bench.py > gpurun_out/final_n1_default.log 2> gpurun_out/final_n1_default.err; echo "bench rc=$?"
tail -c 600 gpurun_out/final_n1_default.log
CMD="python bench.py --steps 2 --warmup 3 --no-microbench --no-cpu-baseline --no-c5 --no-fp32 --no-graph"
timeout 120 $CMD > gpurun_out/r2b_plain.log 2>&1 && \
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/r2b_launches_bench_c3_bf16.csv $CMD > gpurun_out/r2b_ncu_l.log 2>&1
echo "ncu list rc=$?"
timeout 240 ncu --set full --clock-control none --import-source on -k regex:'window_update|emb_fwd_kernel' -s 8 -c 6 -f -o gpurun_out/r2b_kernels $CMD > gpurun_out/r2b_ncu_f.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out/ | tail -6
