timeout 300 python -m pytest tests/test_gpu_parallel.py -x -q 2>&1 | tail -30 > gpurun_out/parallel_n2.log
cat gpurun_out/parallel_n2.log
