#!/bin/bash
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2b_gpu_tests.log 2>&1; tail -5 gpurun_out/r2b_gpu_tests.log
python tools/shard_probe.py 2>&1 | tee gpurun_out/r2_shard_probe.txt
V=v15 bash tools/capture_evidence.sh > gpurun_out/capture_v15.log 2>&1; tail -30 gpurun_out/capture_v15.log
