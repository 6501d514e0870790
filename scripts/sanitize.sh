#!/bin/bash
# compute-sanitizer over a reduced parity subset (one tool per GPU call, see B200_PROFILING.md):
#   TOOL=memcheck|racecheck|initcheck|synccheck bash scripts/sanitize.sh
# Summaries land in gpurun_out/sanitize_$TOOL.log (copied to profiles/ by hand).
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
TOOL=${TOOL:-memcheck}
SUBSET='test_c1_default_scene_fused or test_c4_cornell_depth8 or test_path_chain_bvh_deeper or test_tile_split_pack_unpack_bit_identical[2] or test_multi_buffer_triangle_fetch or test_brute_force_without_bvh or test_wave_engine_sub_frames[3] or test_rebuilt_policy_quantised_nodes_under_transforms[offset] or test_dispatch_with_caller_built'
mkdir -p gpurun_out
# the plain run first: a faulting program under a tool has left the GPU unusable before
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "$SUBSET" > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
tail -1 gpurun_out/sanitize_plain.log
if [ "$TOOL" = "checked" ]; then
  # compute-sanitizer is closed on this pool: the same subset against the -DRT_CHECKED build (device-side bounds asserts)
  RTB200_LIB=$PWD/build/tune/checked.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu -k "not c5_10m" > gpurun_out/sanitize_checked_pytest.log 2>&1
  echo "checked build exit $?"; tail -3 gpurun_out/sanitize_checked_pytest.log; exit 0
fi
EXTRA=""
[ "$TOOL" = "initcheck" ] && EXTRA="--track-unused-memory no"
timeout 3000 compute-sanitizer --tool $TOOL $EXTRA --target-processes all --print-limit 30 --log-file gpurun_out/sanitize_$TOOL.log \
    python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "$SUBSET" > gpurun_out/sanitize_${TOOL}_pytest.log 2>&1
echo "sanitizer $TOOL exit $?"
tail -2 gpurun_out/sanitize_${TOOL}_pytest.log
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|Uninitialized" gpurun_out/sanitize_$TOOL.log | sort | uniq -c | sort -rn | head -20
