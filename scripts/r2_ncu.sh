# one ncu --set full pass over the window kernels (one stream, UNITS-unit launches).  usage: r2_ncu.sh TAG [units] [regex] [skip] [count]
TAG=${1:-r2_ncu}
UNITS=${2:-24}
RX=${3:-k_win_prepass|k_splat_window|k_gather_normalise}
SKIP=${4:-6}
CNT=${5:-3}
O=gpurun_out/$TAG
mkdir -p $O
B="python scripts/ncu_target.py $((2*UNITS)) 3"
export TCW_LANES=1 TCW_CHUNK_UNITS=$UNITS TCW_NO_GRAPH=1
timeout 300 $B > $O/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"$RX" -s $SKIP -c $CNT -o $O/kernels_full $B > $O/ncu.log 2>&1
cat $O/plain.log; tail -3 $O/ncu.log; ls -la $O
