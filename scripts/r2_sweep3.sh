# schedule sweep, second round: usage r2_sweep3.sh TAG  (n_units : lanes chunk ...)
O=gpurun_out/${1:-r2_sweep3}; mkdir -p $O
run() { n=$1; shift; for cfg in "$@"; do set -- $cfg; TCW_LANES=$1 TCW_CHUNK_UNITS=$2 timeout 200 python scripts/ncu_target.py $n 6 window 0 2>&1 | tail -1 | sed "s/^/lanes $1 chunk $2: /" | tee -a $O/sweep.txt; done; }
run 392 "4 14" "4 16" "4 24" "2 32" "2 48" "3 16" "3 24" "2 24" "6 12"
run 196 "4 2" "4 4" "4 7" "2 12" "4 10" "2 24"
run 98 "4 2" "4 3" "4 4" "2 6" "4 6" "2 12"
