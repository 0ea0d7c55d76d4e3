mkdir -p gpurun_out
set -x
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r1_v1.json 2> gpurun_out/bench_r1_v1.err; echo rc=$?
tail -c 3000 gpurun_out/bench_r1_v1.json
for g in 1 2 4 8 16 49; do echo "chunk $g"; TCW_CHUNK_UNITS=$g python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value',round(d['value']),'Mpix/s ms',round(d['ms_per_step'],3),'frac',round(d['roofline']['frac'],4),'step_frac',round(d['roofline']['step_frac'],4),d['roofline']['per_kernel_ms_per_step'],'launches',d['gpu_launches'])
"; done
python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 400 --csv --log-file gpurun_out/launches_r1_v1.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_project_splat -s 30 -c 2 -o gpurun_out/prof_r1_v1_splat python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
