summ() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 ms',round(d['ms_per_step'],3),'iso',{k:round(v,3) for k,v in r['per_kernel_ms_per_step'].items()})"; }
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
timeout 200 python bench.py --steps 30 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | summ "k0pin"
