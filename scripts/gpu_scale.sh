# multi-GPU series on one box: the driver's torchrun command at N = $@ (default 2); weak value, strong, with_gather
O=gpurun_out/scale; mkdir -p $O
for n in ${@:-2}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 10 --warmup 3 > $O/r2_n$n.json 2> $O/r2_n$n.err; echo "N=$n rc=$?"
  tail -3 $O/r2_n$n.err
  python - <<PY
import json
try:
    d=json.loads(open('$O/r2_n$n.json').read().strip().splitlines()[-1])
    print('N=$n weak', round(d['value']), 'ms', round(d['ms_per_step'],3), 'e2e', (d.get('e2e') or {}).get('value'))
    print('  strong', d.get('strong')); print('  with_gather', d.get('with_gather'))
except Exception as e:
    print('no line', e)
PY
done
