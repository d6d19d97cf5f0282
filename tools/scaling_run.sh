#!/bin/bash
# bench.py at N = 1, 2, 4, 8 on one 8-GPU box, plus the multi-GPU tests (run under gpurun --gpus 8)
set -x
V=${V:-v13}
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/tests_${V}_8gpu.log 2>&1; tail -3 gpurun_out/tests_${V}_8gpu.log
python bench.py --steps 40 --warmup 5 > gpurun_out/bench_${V}_n1.json 2> gpurun_out/bench_${V}_n1.err
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
      bench.py --gpus $N --steps 40 --warmup 5 > gpurun_out/bench_${V}_n$N.json 2> gpurun_out/bench_${V}_n$N.err
  echo "N=$N rc=$?"
  tail -1 gpurun_out/bench_${V}_n$N.json | cut -c1-200
done
