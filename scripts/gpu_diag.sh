TAG=${1:-diag}
O=gpurun_out/$TAG; mkdir -p $O
B="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
export TCW_PERSIST=1
timeout 120 $B > $O/plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"k_warp_pipeline2" -s 3 -c 1 -o $O/static $B > $O/ncu.log 2>&1
ls -la $O; tail -2 $O/ncu.log | cut -c1-200
