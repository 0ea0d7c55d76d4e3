timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
for i in 1 2; do timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('value',round(d['value']),'ms',round(d['ms_per_step'],3),'launches',d['gpu_launches'], d['roofline']['step_frac'], d['roofline']['per_kernel_ms_per_step'])"; done
TCW_NO_GRAPH=1 timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('nograph value',round(d['value']),'ms',round(d['ms_per_step'],3))"
