O=gpurun_out/${1:-r2_sweep5}; mkdir -p $O
for off in 0 147 343; do for cfg in "4 2" "4 3" "4 4" "8 2"; do set -- $cfg; TCW_LANES=$1 TCW_CHUNK_UNITS=$2 timeout 200 python scripts/ncu_target.py 49 8 window $off 2>&1 | tail -1 | sed "s/^/first unit $off, lanes $1 chunk $2: /" | tee -a $O/sweep.txt; done; done
