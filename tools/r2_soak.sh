#!/bin/bash
# Stability soak of the shipped build: process-start stress (padded slots, both batch paths) and the tests of the
# concurrent / overlapping / look-ahead / combine paths repeated in fresh processes.
mkdir -p gpurun_out
N=${N:-150} P=6 bash tools/flaky_loop.sh | tail -1
: > gpurun_out/soak_tests.log
for r in $(seq 1 ${REPS:-15}); do
  timeout 600 python -m pytest -q -m gpu -x \
    tests/test_gpu_parity.py::test_back_to_back_launches_with_lookahead_and_overlap \
    tests/test_gpu_parity.py::test_concurrent_host_calls_on_one_handle \
    tests/test_gpu_multi.py::test_server_answer_from_many_threads \
    tests/test_gpu_multi.py::test_combine_slots_single_process \
    tests/test_gpu_multi.py::test_combine_wait_read_reset_frees_the_slot \
    tests/test_gpu_parity.py::test_keyword_prefix_sharing_kernel_matches_oracle \
    tests/test_zz_no_alloc.py 2>&1 | tail -1 >> gpurun_out/soak_tests.log
done
sort gpurun_out/soak_tests.log | sed 's/ in [0-9.]*s//' | uniq -c
