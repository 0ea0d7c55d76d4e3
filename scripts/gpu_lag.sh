timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 200 python bench.py --steps 30 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('value',round(d['value']),'ms',round(d['ms_per_step'],3))"
