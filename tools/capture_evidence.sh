#!/bin/bash
# Evidence for profiles/ of one kernel generation (V names it): GPU tests, bench at two lengths, the reference arm,
# the ncu launch list of the bench command, one `ncu --set full` capture per kernel that matters.
# Usage: V=v14 bash tools/capture_evidence.sh     (1 GPU; every ncu command only after its plain run exited 0)
set -x
V=${V:-v14}
mkdir -p gpurun_out
timeout 900 echo "(GPU tests: see r2e_gpu_tests.log)"
python bench.py > gpurun_out/bench_${V}_default.json 2> gpurun_out/bench_${V}_default.err; echo rc=$?
python bench.py --steps 200 --warmup 5 > gpurun_out/bench_${V}_steps200.json 2> gpurun_out/bench_${V}_steps200.err; echo rc=$?
python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_${V}_ref.json 2> gpurun_out/bench_${V}_ref.err; echo rc=$?
PIRB_BENCH_OTHER=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${V}.csv python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_${V}.log 2>&1; echo rc=$?
python tools/parity_probe.py > gpurun_out/parity_probe_${V}.txt 2>&1; echo rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:answer_kernel -c 1 -o gpurun_out/answer_${V}_even -f python tools/parity_probe.py > gpurun_out/ncu_full_${V}_even.log 2>&1; echo rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:answer_kernel --launch-skip 12 -c 1 -o gpurun_out/answer_${V}_odd -f python tools/parity_probe.py > gpurun_out/ncu_full_${V}_odd.log 2>&1; echo rc=$?
python tools/kernels_probe.py > gpurun_out/kernels_probe_${V}.txt 2>&1; echo rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:keyword_walk_kernel -c 1 -o gpurun_out/keyword_${V} -f python tools/kernels_probe.py > gpurun_out/ncu_full_${V}_keyword.log 2>&1; echo rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:batch_scan_kernel -c 1 -o gpurun_out/batch_scan_${V} -f python tools/kernels_probe.py > gpurun_out/ncu_full_${V}_batch_scan.log 2>&1; echo rc=$?
ls -la gpurun_out/ | tail -20
