# evidence round, part A: tests + bench lines + smoke + ncu launch list (one profiler pass per gpurun call);
# part B (scripts/gpu_round_ncu.sh): ncu --set full captures.  Outputs under gpurun_out/<tag>
TAG=${1:-r1}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi.txt
timeout 900 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; cat $O/bench.json
timeout 300 python bench.py --steps 20 --warmup 3 --deterministic --no-e2e --no-cpu-baseline > $O/bench_det.json 2>> $O/bench.err; cat $O/bench_det.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2>> $O/bench.err; cat $O/bench_ref.json
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
B="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 300 $B > $O/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 800 -c 420 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
ls -la $O
