#!/bin/bash
# round 2, first GPU call: the whole GPU suite, smoke, the bench line (both arms), the process-start stress
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=12 ) > gpurun_out/r2_gpu_tests.log 2>&1; tail -25 gpurun_out/r2_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
( time timeout 900 python bench.py ) > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc=$?"; tail -5 gpurun_out/r2_bench_default.err; cut -c1-1500 gpurun_out/r2_bench_default.json
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 3 ) > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"
N=${N:-300} P=6 bash tools/flaky_loop.sh
