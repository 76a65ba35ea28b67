#!/bin/bash
# compute-sanitizer over one test of every kernel family ("memcheck / racecheck on every kernel in tests").
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck   ->  gpurun_out/r02_sanitizer_<tool>.txt
# One tool per gpurun call (B200_PROFILING.md).  The tests themselves compare every result with the CPU oracle.
TOOL=${1:-memcheck}
OUT=gpurun_out/r02_sanitizer_${TOOL}.txt
mkdir -p gpurun_out
SEL=(
 "tests/test_extractor_gpu.py::test_stages_and_end_to_end[640-480-1000-0]"
 "tests/test_extractor_gpu.py::test_degenerate_inputs"
 "tests/test_frame_gpu.py::test_extractor_epilogue[tum1-5]"
 "tests/test_matcher_gpu.py::test_search_for_initialization[0-0.9-True-100]"
 "tests/test_matcher_gpu.py::test_search_by_projection_local_points[5000-5.0-True]"
 "tests/test_matcher_gpu.py::test_search_by_projection_last_frame[forward-15.0-True-True]"
 "tests/test_matcher_gpu.py::test_search_by_bow[True-4-0.75-True-True]"
 "tests/test_matcher_gpu.py::test_compute_stereo_matches_batch_device"
 "tests/test_matcher_gpu.py::test_distinctive_descriptors"
 "tests/test_guided_gpu.py::test_fuse_search[True-True-3.0-5000]"
 "tests/test_guided_gpu.py::test_search_by_sim3[7.5]"
 "tests/test_guided_gpu.py::test_search_for_triangulation[1-0.1-True-True-True-epipole3]"
 "tests/test_guided_gpu.py::test_search_by_projection_frame_keyframe[10.0-100-True-False]"
 "tests/test_bench_path_gpu.py::test_search_init_consecutive_device_empty_frame_and_two_frames"
 "tests/test_bench_path_gpu.py::test_knn2_device_engines_match_oracle[257-129-popc]"
 "tests/test_bench_path_gpu.py::test_knn2_device_engines_match_oracle[257-129-mma]"
 "tests/test_bench_path_gpu.py::test_knn2_device_engines_match_oracle[4096-4096-mma]"
 "tests/test_bench_path_gpu.py::test_knn2_device_ties_extremes_empty"
)
EXTRA=""
[ "$TOOL" = "racecheck" ] && EXTRA="--racecheck-report all"
{ echo "# compute-sanitizer --tool $TOOL $EXTRA, $(nvidia-smi --query-gpu=name,driver_version --format=csv,noheader), $(date -u +%FT%TZ)"
  echo "# selection: ${SEL[*]}"; } > "$OUT"
timeout ${SANITIZE_TIMEOUT:-1500} compute-sanitizer --tool "$TOOL" $EXTRA --print-limit 50 --error-exitcode 7 \
    python -m pytest -x -q -p no:cacheprovider "${SEL[@]}" >> "$OUT" 2>&1
echo "# exit code $?" >> "$OUT"
grep -E "passed|failed|ERROR SUMMARY|RACECHECK SUMMARY|exit code" "$OUT" | tail -6
