python tools/tf32_err.py
for S in 0 1; do echo "tf32=$S"; B200_INTERACT_TF32=$S timeout 300 python tools/microbench.py --tables 26 --rows 1000 --dtype fp32 --cases 32768:1 --iters 20 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'interact' in d['kernel']: print(d['kernel'], d.get('B'), d['ms'], d.get('GBps'))
"; done
timeout 300 python -m pytest tests -m gpu -x -q -k "interaction or interact" 2>&1 | tail -3
