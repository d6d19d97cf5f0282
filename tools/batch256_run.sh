#!/bin/bash
# batches of fused launches (2^24 x 256 B): two side streams vs one stream, throughput vs latency work split
for Q in 16 64; do
for mode in "A=1" "PIRB_BATCH_LATENCY_SPLIT=1" "PIRB_BATCH_ONE_STREAM=1"; do
  env $mode PIRB_BENCH_QUERIES=$Q python bench.py --steps 5 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$Q $mode', 'GB/s', round(d['value']), 'ms/batch', round(d['ms_per_step'],3), 'q/s', round(d['queries_per_s'],1), 'e2e', round(d['e2e']['value']))"
done; done
