# quick GPU check: tests + short bench of both accumulation modes
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -8
for extra in "" "--deterministic"; do
python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline $extra 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('$extra value',round(d['value']),'Mpix/s ms',round(d['ms_per_step'],3),'frac',round(r['frac'],4),'step_frac',round(r['step_frac'],4),{k:round(v,3) for k,v in r['per_kernel_ms_per_step'].items()},'launches',d['gpu_launches'])
    else: print(l.rstrip())
"
done
