# wider schedule sweep of the default path: lanes x units per chunk, for a 392-unit step and a 49-unit clip
TAG=${1:-r2_sweep2}
O=gpurun_out/$TAG
mkdir -p $O
for n in 392 49; do
for cfg in "4 2" "1 24" "2 8" "2 12" "2 16" "2 24" "3 8" "4 4" "4 6" "4 8" "4 12" "6 4" "6 8" "8 3" "8 6"; do
  set -- $cfg
  TCW_LANES=$1 TCW_CHUNK_UNITS=$2 timeout 200 python scripts/ncu_target.py $n 6 window 0 2>&1 | tail -1 | sed "s/^/lanes $1 chunk $2: /" | tee -a $O/sweep.txt
done
done
