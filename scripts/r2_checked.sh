# memory-safety evidence without compute-sanitizer (closed on this pool): build the library with -DTCW_BOUNDS_CHECK=1 (every
# window / slab / pool access of the default path range-checked, violations skipped and flagged in status bit 8), then run
# the small fixed program and the GPU parity tests on that build.  A violation raises in Warper.check_status.
O=gpurun_out/r2_checked
mkdir -p $O
( cd trajectorycrafter_b200 && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
    -shared -Xcompiler -fPIC,-fvisibility=hidden -DTCW_BOUNDS_CHECK=1 -I ../include -o libtcwarp_chk.so csrc/tcwarp.cu ) || exit 1
export TCW_LIB=$PWD/trajectorycrafter_b200/libtcwarp_chk.so
timeout 300 python scripts/sanitize_target.py > $O/target.log 2>&1; echo "target rc=$?"; tail -3 $O/target.log
timeout 900 python -m pytest tests -x -q -m gpu > $O/pytest_gpu_checked.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu_checked.log
