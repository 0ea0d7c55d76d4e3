# First GPU call of the next round: A/B of the prepared compile-time experiments against the default library.
# Builds the variants next to libtcwarp.so (they travel with the gpurun snapshot when built before the call):
#   bash scripts/next_round.sh build      (here, CPU only)
#   gpurun --timeout 1500 -- 'bash scripts/next_round.sh run'
# For every variant: the GPU parity suite on that library (TCW_LIB), then bench.py without e2e / CPU legs.
P=trajectorycrafter_b200
NVCC="nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -shared -Xcompiler -fPIC,-fvisibility=hidden -I include"
VARIANTS="rows2:-DTCW_ROWS2=1 rows2_b3:-DTCW_ROWS2=1,-DTCW_SPLAT_MINB=3 cm4:-DTCW_CONST_M=1 cm5:-DTCW_CONST_M=1,-DTCW_SPLAT_MINB=5"
if [ "$1" = build ]; then
  for v in $VARIANTS; do n=${v%%:*}; f=$(echo ${v#*:} | tr ',' ' '); echo "== $n ($f)"; $NVCC $f -Xptxas=-v -o $P/libtcwarp_$n.so $P/csrc/tcwarp.cu 2>&1 | grep -A2 "k_splat_pair8PairArgsi\|k_splat_pair_cm" | grep "registers\|spill"; done
  (cd bench_micro && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o micro_red_patterns micro_red_patterns.cu)
  exit 0
fi
O=gpurun_out/next; mkdir -p $O
summ() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$1 ms',round(d['ms_per_step'],3),'Gpix/s',round(d['value']/1e3,1),'K2 alone',round(r['per_kernel_ms_per_step']['project_splat'],3))"; }
timeout 120 bench_micro/micro_red_patterns | tee $O/micro_red_patterns.txt
timeout 200 python bench.py --steps 30 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | summ default
for v in $VARIANTS; do n=${v%%:*}; L=$PWD/$P/libtcwarp_$n.so; [ -f $L ] || continue
  TCW_LIB=$L timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -m gpu 2>&1 | tail -1
  TCW_LIB=$L timeout 200 python bench.py --steps 30 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | summ $n
done
