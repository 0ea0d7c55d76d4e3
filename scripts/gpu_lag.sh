timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 3 2>/dev/null | tee gpurun_out/bench_lanes.json | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value',round(d['value']),'ms',round(d['ms_per_step'],3),'e2e',d['e2e'],'cpu',d['cpu_baseline'],'launches',d['gpu_launches'])"
TCW_PERSIST=1 timeout 120 python -m pytest tests -x -q -m gpu 2>&1 | tail -1
TCW_V3=1 timeout 120 python -m pytest tests -x -q -m gpu 2>&1 | tail -1
