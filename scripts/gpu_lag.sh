timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/bench_pack.json; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_pack.json').read().strip().splitlines()[-1])
print('value',round(d['value']),'ms',round(d['ms_per_step'],3),'frac',round(d['roofline']['frac'],3),'step_frac',round(d['roofline']['step_frac'],3))
print('pack',d.get('pack_epilogue'))
PY
