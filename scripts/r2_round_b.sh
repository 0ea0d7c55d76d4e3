# round-2 evidence, part B: ncu passes on the small fixed program (scripts/ncu_target.py), each after its plain run exited 0.
#   1. launch list of one whole 392-unit step in the timed schedule's launch shape (4 lanes x 12 units: 33 chunks x 3 kernels),
#      caches left alone (durations + DRAM bytes)
#   2. ncu --set full of the three kernels in the launch shape the timed schedule runs (12 units per launch)
#   3. the same at 24 units per launch (the shape roofline.kernel_alone is timed on)
O=gpurun_out/${1:-r2_round_b}
mkdir -p $O
export TCW_NO_GRAPH=1
B="python scripts/ncu_target.py 392 3 window 0"
timeout 300 $B > $O/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -s 198 -c 99 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
export TCW_LANES=1 TCW_CHUNK_UNITS=12
B="python scripts/ncu_target.py 48 3 window 160"
timeout 300 $B > $O/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"k_win_prepass|k_splat_window|k_gather_normalise" -s 15 -c 3 -o $O/kernels_full_12units $B > $O/ncu2.log 2>&1
export TCW_CHUNK_UNITS=24
timeout 300 $B > $O/plain3.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"k_win_prepass|k_splat_window|k_gather_normalise" -s 9 -c 3 -o $O/kernels_full_24units $B > $O/ncu3.log 2>&1
for f in plain plain2 plain3 ncu2 ncu3; do tail -n 2 $O/$f.log; done; ls -la $O
