# tests + bench + ncu launch list + one full capture of the dominant kernel; outputs under gpurun_out/<tag>
TAG=${1:-r1}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi.txt
python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
python bench.py --steps 20 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; cat $O/bench.json
python bench.py --steps 20 --warmup 3 --deterministic --no-e2e --no-cpu-baseline > $O/bench_det.json 2>> $O/bench.err; cat $O/bench_det.json
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2>> $O/bench.err; cat $O/bench_ref.json
python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > $O/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > $O/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_project_splat -s 6 -c 1 -o $O/splat_full python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > $O/ncu2.log 2>&1
ls -la $O
