# multi-GPU weak scaling on one 8-GPU box: the driver's bench command at N=8 (full line incl. e2e) and N=2/4 without e2e
O=gpurun_out/scale; mkdir -p $O
for n in 8 4 2; do
  extra=""; [ $n != 8 ] && extra="--no-e2e"
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 3 $extra > $O/final_n$n.json 2> $O/final_n$n.err; echo "N=$n rc=$?"
  python - <<PY
import json
d=json.loads(open('$O/final_n$n.json').read().strip().splitlines()[-1])
print('N=$n value', round(d['value']), 'ms', round(d['ms_per_step'],3), 'e2e', (d.get('e2e') or {}).get('value'), 'per_rank', d.get('per_rank',{}).get('ms_per_step'), 'gather', d.get('gather'))
PY
done
