#!/bin/bash
# 2-GPU check: multi-GPU tests + bench at N=2, with the in-kernel arrival signal and with the separate signalling kernel
set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for mode in fold kernel; do
  if [ $mode = kernel ]; then export PIRB_SIGNAL_KERNEL=1; else unset PIRB_SIGNAL_KERNEL; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/bench_n2_$mode.json 2> gpurun_out/bench_n2_$mode.err
  echo "$mode rc=$?"; tail -1 gpurun_out/bench_n2_$mode.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel_ms'], d['gpu_launches'], d['clocks'])"
done
