#!/bin/bash
# compute-sanitizer over the parity tests (run on the GPU box through gpurun; logs under gpurun_out/)
O=gpurun_out
CS="compute-sanitizer --error-exitcode 9 --launch-timeout 0"
run() { # name tool tests...
  n=$1; t=$2; shift 2
  timeout 900 $CS --tool $t python -m pytest "$@" -q -x -p no:cacheprovider > $O/r2_sanitizer_${n}.log 2>&1
  echo "$n rc=$? $(grep -c 'ERROR SUMMARY' $O/r2_sanitizer_${n}.log) $(grep 'ERROR SUMMARY\|passed\|failed' $O/r2_sanitizer_${n}.log | tr '\n' ' ' | cut -c1-300)"
}
run memcheck_scan memcheck tests/test_gpu_scan.py tests/test_zz_no_alloc.py
run memcheck_actor memcheck tests/test_gpu_actor.py -k "golden or ragged"
run memcheck_learner memcheck tests/test_gpu_learner.py -k "golden or many_tiles or pickle"
run racecheck_scan racecheck tests/test_gpu_scan.py -k "golden or both_kernels"
run racecheck_actor racecheck tests/test_gpu_actor.py -k "golden or ragged"
run racecheck_learner racecheck tests/test_gpu_learner.py -k "epoch_golden"
run synccheck_all synccheck tests/test_gpu_scan.py tests/test_gpu_actor.py -k "golden or ragged"
