# quick GPU check: tests + short bench (v4 pipeline, v3 pipeline for A/B)
TAG=${1:-quick}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 | tee $O/pytest_gpu.log
summ() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('$1 value',round(d['value']),'Mpix/s ms',round(d['ms_per_step'],3),'frac',round(r['frac'],4),'step_frac',round(r['step_frac'],4),{k:round(v,3) for k,v in r['per_kernel_ms_per_step'].items()},'launches',d['gpu_launches'])
    else: print(l.rstrip())
"; }
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | tee $O/bench_v4.json | summ "v4"
TCW_V3=1 timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | summ "v3"
