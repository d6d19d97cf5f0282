#!/bin/bash
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -q ) > gpurun_out/r2e_gpu_tests.log 2>&1; tail -6 gpurun_out/r2e_gpu_tests.log; grep -E "^FAILED|^E  " gpurun_out/r2e_gpu_tests.log | head -20
V=v16 bash tools/capture_evidence.sh > gpurun_out/capture_v16.log 2>&1; grep -E "rc=|passed|failed" gpurun_out/capture_v16.log | head -30
PIRB_PHASE_CASES="1:auto,8:auto" bash tools/phase_debug.sh > gpurun_out/phase_v16.txt 2>&1; tail -24 gpurun_out/phase_v16.txt
python tools/shard_probe.py 2>&1 | tee gpurun_out/r2_shard_probe_v16.txt
