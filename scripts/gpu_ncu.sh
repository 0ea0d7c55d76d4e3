# one full ncu capture of kernels matching $2 (regex), skipping $3 launches, capturing $4; tag $1; rest = bench args
# caches are NOT flushed between kernels (--cache-control none) so L2 residency is the pipeline's own
TAG=$1; PAT=$2; SKIP=${3:-10}; CNT=${4:-2}; shift 4
O=gpurun_out/$TAG; mkdir -p $O
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline $@"
$CMD > $O/plain.log 2>&1 && \
ncu --set full --clock-control none --cache-control none --import-source on -k regex:$PAT -s $SKIP -c $CNT -o $O/prof $CMD > $O/ncu.log 2>&1
tail -3 $O/ncu.log; ls -la $O
