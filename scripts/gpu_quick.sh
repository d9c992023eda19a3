#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/q
mkdir -p $O
timeout 600 python -m pytest tests/test_model_gpu.py -q -m gpu --timeout=600 2>&1 | tail -2
T="tests/test_zz_no_alloc.py tests/test_moment_spread_gpu.py::test_fused_generator_and_rhs_pass_matches_exact_nudft tests/test_moment_spread_gpu.py::test_moment_spread_batched_rows_are_independent tests/test_misc_gpu.py::test_convolution_cg_matches_fft_cg"
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest $T -x -q -m gpu --timeout=1100 > $O/memcheck.log 2>&1
echo "memcheck rc=$?"; grep -E "ERROR SUMMARY|passed|failed|Invalid|out of bounds" $O/memcheck.log | head -8
timeout 1500 compute-sanitizer --tool racecheck --error-exitcode 7 python -m pytest tests/test_moment_spread_gpu.py::test_fused_generator_and_rhs_pass_matches_exact_nudft tests/test_zz_no_alloc.py::test_static_scratch_bbox_back_to_back_without_host_sync -x -q -m gpu --timeout=1400 > $O/racecheck.log 2>&1
echo "racecheck rc=$?"; grep -E "RACECHECK SUMMARY|passed|failed|hazard" $O/racecheck.log | head -8
