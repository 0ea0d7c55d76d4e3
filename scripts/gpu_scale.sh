# multi-GPU weak-scaling check on one box: bench at N=8 (full line incl. e2e) and N=4; outputs under gpurun_out/scale
O=gpurun_out/scale; mkdir -p $O
nvidia-smi -L | wc -l
for n in 8 4; do
  extra=""; [ $n = 4 ] && extra="--no-e2e"
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 3 $extra > $O/bench_n$n.json 2> $O/bench_n$n.err; echo "N=$n rc=$?"; tail -c 3000 $O/bench_n$n.json
done
