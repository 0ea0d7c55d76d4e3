summ() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 ms',round(d['ms_per_step'],3))"; }
for p in 0 1 0 1; do TCW_PRIO=$p timeout 200 python bench.py --steps 30 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | summ "prio=$p"; done
TCW_PRIO=1 TCW_NO_GRAPH=1 timeout 200 python bench.py --steps 30 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | summ "prio=1 nograph"
