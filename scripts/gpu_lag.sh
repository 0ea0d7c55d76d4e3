timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
timeout 200 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --deterministic 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('det value',round(d['value']),'ms',round(d['ms_per_step'],3),'launches',d['gpu_launches'], d['roofline']['per_kernel_ms_per_step'])"
timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['roofline']; print('value',round(d['value']),'ms',round(d['ms_per_step'],3),'frac',r['frac'],'iso',r['isolated'])"
