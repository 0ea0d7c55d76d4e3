# round-2 quick GPU check: GPU tests + short bench of the default (deterministic window) path and the opt-in float-reduction path
TAG=${1:-r2_quick}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi.txt
timeout 900 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; cat $O/bench.json; tail -5 $O/bench.err
timeout 300 python bench.py --steps 20 --warmup 3 --accumulation atomic --no-e2e --no-cpu-baseline > $O/bench_atomic.json 2>> $O/bench.err; cat $O/bench_atomic.json
