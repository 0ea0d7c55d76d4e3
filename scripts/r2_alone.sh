# per-kernel "alone" times (bench.py profile pass) for every build of the library found
for lib in trajectorycrafter_b200/libtcwarp*.so; do
  TCW_LIB=$PWD/$lib timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']
        print('$(basename $lib)', 'step', round(d['ms_per_step'],3), 'alone', {k:round(v['ms_per_step'],3) for k,v in r['kernel_alone'].items()}, 'in_schedule', {k:round(v,2) for k,v in r['in_schedule']['per_kernel_ms_per_step'].items()})
"
done
