# sweep of measurement knobs (environment settings) on the full 392-unit step, device-timed, default schedule
# usage: r2_env_sweep.sh TAG 'ENV=.. ENV=..' 'ENV=..' ...
TAG=${1:-r2_env}; shift
O=gpurun_out/$TAG
mkdir -p $O
for cfg in "$@"; do
  for rep in 1 2; do
    env $cfg timeout 200 python scripts/ncu_target.py 392 6 window 0 2>&1 | tail -1 | sed "s|^|[$cfg] |" | tee -a $O/sweep.txt
  done
done
