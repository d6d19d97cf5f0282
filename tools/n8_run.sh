#!/bin/bash
# 8-GPU box: combine variants on the headline workload, then BASELINE configs[4] (2^28 x 32 B, index and keyword
# queries) at 1 and 8 GPUs
set -x
run() {  # name, nproc, extra env...
  name=$1; np=$2; shift 2
  if [ $np = 1 ]; then
    env "$@" python bench.py --steps 20 --warmup 5 > gpurun_out/$name.json 2> gpurun_out/$name.err
  else
    env "$@" timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) \
      bench.py --gpus $np --steps 20 --warmup 5 > gpurun_out/$name.json 2> gpurun_out/$name.err
  fi
  echo "$name rc=$?"; tail -1 gpurun_out/$name.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel_ms'], d['gpu_launches'])"
}
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -2
run c3_n8_local 8 A=1
run c3_n8_direct 8 PIRB_COMBINE_DIRECT=1
run c3_n8_local_b 8 A=1
run c3_n8_sigkernel 8 PIRB_SIGNAL_KERNEL=1
run c5_index_n8 8 PIRB_BENCH_LOG_ROWS=28 PIRB_BENCH_SLOT_BYTES=32
run c5_keyword_n8 8 PIRB_BENCH_LOG_ROWS=28 PIRB_BENCH_SLOT_BYTES=32 PIRB_BENCH_KEYWORD=1
run c5_index_n1 1 PIRB_BENCH_LOG_ROWS=28 PIRB_BENCH_SLOT_BYTES=32
run c5_keyword_n1 1 PIRB_BENCH_LOG_ROWS=28 PIRB_BENCH_SLOT_BYTES=32 PIRB_BENCH_KEYWORD=1
