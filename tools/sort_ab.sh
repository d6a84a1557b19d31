python -m pytest tests -m gpu -x -q -k "sort or dedup or bwd or no_alloc or fullsize" 2>&1 | tail -2
for S in 0 1; do echo "onesweep=$S"; B200_SORT_ONESWEEP=$S python tools/microbench.py --c3 --no-interact --dtype fp32 --act bf16 --cases 32768:1 --iters 20 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'bwd' in d['kernel']: print(d['kernel'], d['B'], d['P'], d['ms'], d.get('GBps'))
"; B200_SORT_ONESWEEP=$S python tools/microbench.py --no-interact --dtype fp32 --cases 262144:1,8192:32 --iters 10 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'bwd' in d['kernel']: print(d['kernel'], d['B'], d['P'], d['ms'], d.get('GBps'))
"; done
