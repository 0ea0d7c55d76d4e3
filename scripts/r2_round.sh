# round-2 evidence, part A (one profiler pass per gpurun call): GPU tests, the driver's bench commands, smoke, then the ncu
# launch list of one step (durations + DRAM bytes per launch).  Part B = scripts/r2_ncu.sh (ncu --set full captures).
TAG=${1:-r2_final}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi.txt
timeout 900 python -m pytest tests -x -q -m gpu -s > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench.err; echo "ref rc=$?"; cut -c1-300 $O/bench_ref.json
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench.json 2>> $O/bench.err; echo "bench rc=$?"; cut -c1-400 $O/bench.json
timeout 300 python bench.py --steps 20 --warmup 5 --accumulation atomic --no-e2e --no-cpu-baseline --no-extras > $O/bench_atomic.json 2>> $O/bench.err; cut -c1-300 $O/bench_atomic.json
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
export TCW_NO_GRAPH=1
B="python scripts/ncu_target.py 392 3 window 0"
timeout 300 $B > $O/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 198 -c 99 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
tail -2 $O/plain.log; ls -la $O
