#!/bin/bash
# GPU-box job 6 (round 2): compute-sanitizer racecheck / memcheck on the
# shared-memory accumulation paths (scatter tiles, split heavy tiles, sparse
# coverage) and the line FFTs; summaries to profiles/.
O=gpurun_out
mkdir -p $O
SEL="c1_reference_test_generator_fp32 or split_heavy_tiles or sparse_coverage or edge_cases or results_are_deterministic"
( time timeout 1500 compute-sanitizer --tool racecheck --print-limit 20 python -m pytest tests/test_gpu_parity.py -q -x -k "$SEL" ) > $O/r02_sanitizer_racecheck.log 2>&1
grep -E "RACECHECK SUMMARY|passed|failed|error" $O/r02_sanitizer_racecheck.log | tail -5
( time timeout 1500 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_gpu_parity.py -q -x -k "$SEL or golden_vectors or c2_uniform_box_scaled or c3_mid_like" ) > $O/r02_sanitizer_memcheck.log 2>&1
grep -E "ERROR SUMMARY|passed|failed" $O/r02_sanitizer_memcheck.log | tail -5
( time timeout 900 compute-sanitizer --tool synccheck --print-limit 20 python -m pytest tests/test_gpu_parity.py -q -x -k "c1_reference_test_generator_fp32 or split_heavy_tiles" ) > $O/r02_sanitizer_synccheck.log 2>&1
grep -E "ERROR SUMMARY|passed|failed" $O/r02_sanitizer_synccheck.log | tail -5
