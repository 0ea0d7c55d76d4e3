summ() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 value',round(d['value']),'ms',round(d['ms_per_step'],3),'frac',round(r['frac'],3),'iso',{k:round(v,3) for k,v in r['per_kernel_ms_per_step'].items()})"; }
for v in "" _contig; do TCW_LIB=$PWD/trajectorycrafter_b200/libtcwarp$v.so timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | summ "lib$v"; done
