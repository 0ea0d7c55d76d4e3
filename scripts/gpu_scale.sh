O=gpurun_out/scale; mkdir -p $O
timeout 300 python -m pytest tests/test_parity_gpu.py -x -q -m gpu -k "graph or schedule or per_unit" 2>&1 | tail -2
timeout 300 python scripts/step_trace.py 12 2>&1 | grep "^rank" | tee $O/trace_n1_res.txt
timeout 300 python bench.py --warmup 3 --steps 20 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('n1 value',round(d['value']),'ms',round(d['ms_per_step'],3),'host',d['config']['host_enqueue_ms_per_step'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 scripts/step_trace.py 12 2>&1 | grep "^rank" | tee $O/trace_n8_res.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --warmup 3 --steps 20 --no-e2e > $O/n8_res.json 2> $O/n8_res.err; python - <<PY
import json
d=json.loads(open('$O/n8_res.json').read().strip().splitlines()[-1])
print('n8 resident', 'value', round(d['value']), 'ms', round(d['ms_per_step'],3), 'host', d['config']['host_enqueue_ms_per_step'], 'per_rank', d.get('per_rank'))
PY
