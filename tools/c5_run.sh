#!/bin/bash
# BASELINE configs[4] on ONE GPU: 2^28 x 32 B (8 GiB), index and keyword queries; and 2^26 x 64 B (binary split, many tasks per CTA)
mkdir -p gpurun_out
for cfg in "c5_index PIRB_BENCH_LOG_ROWS=28 PIRB_BENCH_SLOT_BYTES=32" "c5_index_nowalker PIRB_BENCH_LOG_ROWS=28 PIRB_BENCH_SLOT_BYTES=32 PIRB_TOP_WARP=0" "c5_keyword PIRB_BENCH_LOG_ROWS=28 PIRB_BENCH_SLOT_BYTES=32 PIRB_BENCH_KEYWORD=1"; do
  set -- $cfg; name=$1; shift
  env "$@" timeout 600 python bench.py --steps ${STEPS:-10} --warmup 3 > gpurun_out/${name}_${V:-v13}.json 2> gpurun_out/${name}_${V:-v13}.err
  echo "$name rc=$?"; tail -1 gpurun_out/${name}_${V:-v13}.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value'],1), 'GB/s', round(d['ms_per_step'],3), 'ms e2e', round(d['e2e']['value'],1))"
done
