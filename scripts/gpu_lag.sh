summ() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 ms',round(d['ms_per_step'],3),'K2 iso',round(r['per_kernel_ms_per_step']['project_splat'],3))"; }
TCW_LIB=$PWD/trajectorycrafter_b200/libtcwarp_m2.so timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for v in _m1 _m2 _m2b3; do TCW_LIB=$PWD/trajectorycrafter_b200/libtcwarp$v.so timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | summ "lib$v"; done
