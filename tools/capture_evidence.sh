#!/bin/bash
timeout 900 ncu --set full --clock-control none --import-source on -k regex:answer_kernel -c 1 -o gpurun_out/answer_${V}_even -f python tools/parity_probe.py > gpurun_out/ncu_full_${V}_even.log 2>&1; echo rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:answer_kernel --launch-skip 12 -c 1 -o gpurun_out/answer_${V}_odd -f python tools/parity_probe.py > gpurun_out/ncu_full_${V}_odd.log 2>&1; echo rc=$?
# of the answer kernel, and the in-kernel phase counters.  V names the output generation.
set -x
V=${V:-v13}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_${V}.log 2>&1; tail -3 gpurun_out/gpu_tests_${V}.log
python bench.py > gpurun_out/bench_${V}_default.json 2> gpurun_out/bench_${V}_default.err; echo rc=$?
python bench.py --steps 40 --warmup 5 > gpurun_out/bench_${V}.json 2> gpurun_out/bench_${V}.err; echo rc=$?
python bench.py --steps 400 --warmup 5 > gpurun_out/bench_${V}_long.json 2> gpurun_out/bench_${V}_long.err; echo rc=$?
python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_${V}_ref.json 2> gpurun_out/bench_${V}_ref.err; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${V}.csv python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_${V}.log 2>&1; echo rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:answer_kernel -c 2 -o gpurun_out/answer_${V} -f python tools/parity_probe.py > gpurun_out/ncu_full_${V}.log 2>&1; echo rc=$?
bash tools/phase_debug.sh > gpurun_out/phase_${V}.log 2>&1; echo rc=$?
ls -la gpurun_out/ | tail -12
