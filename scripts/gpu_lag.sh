summ() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print('$1 value',round(d['value']),'Mpix/s ms',round(d['ms_per_step'],3),{k:round(v,3) for k,v in r['per_kernel_ms_per_step'].items()},'launches',d['gpu_launches'])
"; }
for c in 6 8 10 12 16; do TCW_CHUNK_UNITS=$c timeout 90 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | summ "H8 dual chunk$c"; done
for c in 8 12 24; do TCW_SERIAL=1 TCW_CHUNK_UNITS=$c timeout 90 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | summ "H8 serial chunk$c"; done
