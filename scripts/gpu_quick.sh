# quick GPU check: tests + short bench (fp32 pipeline, fp32 per-chunk launches, deterministic)
TAG=${1:-quick}
O=gpurun_out/$TAG
mkdir -p $O
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -8 | tee $O/pytest_gpu.log
summ() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('$1 value',round(d['value']),'Mpix/s ms',round(d['ms_per_step'],3),'frac',round(r['frac'],4),'step_frac',round(r['step_frac'],4),{k:round(v,3) for k,v in r['per_kernel_ms_per_step'].items()},'launches',d['gpu_launches'])
    else: print(l.rstrip())
"; }
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | summ "pipeline"
TCW_NO_PIPELINE=1 TCW_CHUNK_UNITS=8 timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | summ "chunk8"
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --deterministic 2>&1 | summ "det"
