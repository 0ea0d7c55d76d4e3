# evidence round, part B: one ncu --set full pass over the dominant kernels at the launch shape the roofline figure
# is quoted on (one stream, 24-unit launches); run only after part A's plain bench exited 0
TAG=${1:-r1}
O=gpurun_out/$TAG
mkdir -p $O
B="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
export TCW_SERIAL=1 TCW_CHUNK_UNITS=24
timeout 300 $B > $O/plain24.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"k_splat_pair|k_normalise_pair|k_zmax_quad|k_resize_pack" -s 11 -c 4 -o $O/kernels_full_24units $B > $O/ncu3.log 2>&1
tail -3 $O/ncu3.log; ls -la $O
