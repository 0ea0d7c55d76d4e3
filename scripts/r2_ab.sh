# A/B of two builds of the library on the full 392-unit step (device-timed, default schedule) + the new GPU tests
TAG=${1:-r2_ab}
O=gpurun_out/$TAG
mkdir -p $O
for lib in libtcwarp.so libtcwarp_alt.so; do
  [ -f trajectorycrafter_b200/$lib ] || continue
  TCW_LIB=$PWD/trajectorycrafter_b200/$lib timeout 200 python scripts/ncu_target.py 392 6 window 0 2>&1 | tail -2 | sed "s/^/$lib: /" | tee -a $O/ab.txt
done
timeout 900 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
