# schedule sweep of the default path: lanes x units per chunk (device-timed step, no profiling passes)
TAG=${1:-r2_sweep}
O=gpurun_out/$TAG
mkdir -p $O
for cfg in "4 2" "4 3" "6 1" "6 2" "8 1" "8 2" "8 3" "4 6" "2 12"; do
  set -- $cfg
  TCW_LANES=$1 TCW_CHUNK_UNITS=$2 timeout 200 python scripts/ncu_target.py 392 6 window 0 2>&1 | tail -1 | sed "s/^/lanes $1 chunk $2: /" | tee -a $O/sweep.txt
done
