# NOTE (round 2): compute-sanitizer is CLOSED on this pool (profiles/r2m_sanitizer_closed.txt is what it prints); the bounds
# evidence of this repository is tools/check_bounds_gpu.sh (the GPU suite against a -DORB_CHECKS build).  Kept for pools where
# the tool runs.
# usage: tools/sanitize_gpu.sh [label]  -- compute-sanitizer over the kernels with shared cursors, warp-aggregated atomics,
# shared-memory stacks and cross-stream flags: memcheck (out-of-bounds / misaligned accesses, leaks of
# device memory at exit are not the point), racecheck (shared-memory hazards: the HybridStack columns, k_ltc's staging)
# and synccheck, each over a small-scene subset of the GPU tests.  Summaries land in gpurun_out/<label>_sanitize_*.log.
L=${1:-r2}
mkdir -p gpurun_out
SEL="test_frames_vs_oracle or test_refit_after_moving_meshes or test_band_handoff_through_peer_stores or test_two_frame_slots_equal_one or test_row_tiles_are_bit_identical or test_path_persistent_equals_wavefront or test_quantised_node_stream_kernel"
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 99 --print-limit 20 \
     python -m pytest tests/test_gpu_frames.py -q -x -k "$SEL" > gpurun_out/${L}_sanitize_$tool.log 2>&1
  echo "$tool rc=$?"
  grep -E "ERROR SUMMARY|passed|failed|RACECHECK SUMMARY|Hazard" gpurun_out/${L}_sanitize_$tool.log | tail -5
done
