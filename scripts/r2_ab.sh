# A/B of every build of the library found next to the default one, on the full 392-unit step (device-timed, default schedule)
TAG=${1:-r2_ab}
O=gpurun_out/$TAG
mkdir -p $O
for lib in trajectorycrafter_b200/libtcwarp*.so; do
  TCW_LIB=$PWD/$lib timeout 200 python scripts/ncu_target.py 392 6 window 0 2>&1 | tail -1 | sed "s|^|$(basename $lib): |" | tee -a $O/ab.txt
done
