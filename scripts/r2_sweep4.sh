O=gpurun_out/${1:-r2_sweep4}; mkdir -p $O
for cfg in "4 12" "3 16" "6 8" "4 16" "3 24"; do set -- $cfg; TCW_LANES=$1 TCW_CHUNK_UNITS=$2 timeout 200 python scripts/ncu_target.py 392 6 window 0 2>&1 | tail -1 | sed "s/^/lanes $1 chunk $2: /" | tee -a $O/sweep.txt; done
for w in 8 4; do for cfg in "4 2" "4 3" "4 4" "6 2" "8 1" "8 2" "3 4" "2 6"; do set -- $cfg; TCW_LANES=$1 TCW_CHUNK_UNITS=$2 timeout 200 python scripts/mixed49.py $w 3 2>&1 | tail -1 | sed "s/^/lanes $1 chunk $2: /" | tee -a $O/sweep.txt; done; done
