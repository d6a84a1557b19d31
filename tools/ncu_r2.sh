set -x
CMD="python bench.py --steps 2 --warmup 3 --no-microbench --no-cpu-baseline --no-c5 --no-graph"
$CMD > gpurun_out/r2_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/r2a_launches_bench_c3_bf16.csv $CMD > gpurun_out/r2_ncu_l.log 2>&1
echo "rc=$?"
$CMD > gpurun_out/r2_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'window_update|emb_fwd_kernel|hot_join|window_join' -s 8 -c 8 -f -o gpurun_out/r2_kernels $CMD > gpurun_out/r2_ncu_f.log 2>&1
echo "rc=$?"
ls -la gpurun_out/ | tail -5
