# sweep an env var over values: usage gpu_sweep.sh VAR "v1 v2 ..." [bench args]
VAR=$1; VALS=$2; shift 2
summ() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('$1 value',round(d['value']),'Mpix/s ms',round(d['ms_per_step'],3),'frac',round(r['frac'],4),'step_frac',round(r['step_frac'],4),{k:round(v,3) for k,v in r['per_kernel_ms_per_step'].items()},'launches',d['gpu_launches'])
    else: print(l.rstrip())
"; }
for v in $VALS; do env $VAR=$v timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline "$@" 2>&1 | summ "$VAR=$v"; done
