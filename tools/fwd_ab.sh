python -m pytest tests/test_gpu_ops.py tests/test_gpu_fullsize.py -x -q -k "fwd or full" 2>&1 | tail -2
for P in 0 1; do for C in 3 4; do echo "persist=$P ctas=$C"; B200_FWD_PERSIST=$P B200_FWD_CTAS=$C python tools/microbench.py --c3 --no-interact --dtype fp32 --act bf16 --cases 32768:1 --iters 20 2>/dev/null | grep '"emb_fwd"' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['kernel'], d['B'], d['P'], d['ms'], d['GBps'])
"; [ $P = 0 ] && break; done; done
for P in 1; do echo "uniform persist=$P"; B200_FWD_PERSIST=$P python tools/microbench.py --no-interact --dtype fp32 --cases 262144:1,8192:32,2048:128 --iters 10 2>/dev/null | grep '"emb_fwd"' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['kernel'], d['B'], d['P'], d['ms'], d['GBps'])
"; done
