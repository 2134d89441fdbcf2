# usage: tools/check_bounds_gpu.sh [label]
# The bounds / index evidence of this repo now that compute-sanitizer is closed on the pool
# (profiles/r2m_sanitizer_closed.txt): liborb200_chk.so is the product library compiled with -DORB_CHECKS --
# device-side assertions on claimed work-item indices, hit-record slots, stack depths, node / triangle references,
# queue appends and filter tap coordinates (vec.cuh ORB_DEVICE_ASSERT) -- and the GPU suite runs against it through
# the same C-ABI (ORB200_LIB).  A violated assertion prints file:line and fails the test that triggered it.
# Build here (no GPU needed):  tools/check_bounds_gpu.sh build
L=${1:-r2}
CHK=optix_renderer_b200/liborb200_chk.so
STALE=$(find optix_renderer_b200/csrc include -newer $CHK 2>/dev/null | head -1)
if [ "$L" = build ] || [ ! -f $CHK ] || [ -n "$STALE" ]; then
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -DORB_CHECKS \
       -o $CHK optix_renderer_b200/csrc/orb200.cu || exit 1
  [ "$L" = build ] && exit 0
fi
mkdir -p gpurun_out
ORB200_LIB=$PWD/$CHK timeout 1700 python -m pytest tests/test_gpu_frames.py tests/test_gpu_units.py tests/test_zz_no_alloc.py \
   -m gpu -q -x 2>&1 | tail -15 > gpurun_out/${L}_checked_build.log
echo "assertions compiled in: $(cuobjdump -elf $CHK 2>/dev/null | grep -c '__assertfail')" >> gpurun_out/${L}_checked_build.log
tail -4 gpurun_out/${L}_checked_build.log
