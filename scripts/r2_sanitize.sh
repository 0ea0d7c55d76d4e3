# compute-sanitizer on the small fixed program: ONE tool per gpurun call (B200_PROFILING.md).  usage: r2_sanitize.sh memcheck|racecheck
TOOL=${1:-memcheck}
O=gpurun_out/r2_sanitize
mkdir -p $O
timeout 300 python scripts/sanitize_target.py > $O/plain_$TOOL.log 2>&1 && \
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python scripts/sanitize_target.py > $O/$TOOL.log 2>&1
echo "rc=$?"; tail -5 $O/plain_$TOOL.log; tail -12 $O/$TOOL.log
