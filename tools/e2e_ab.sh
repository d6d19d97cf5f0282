# host call fast path (last CTA writes the pinned buffer) against memset + kernel + copy (development aid)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/e2e_ab.log
for mode in fast copy; do
  if [ $mode = copy ]; then export PIRB_HOST_COPY_PATH=1; else unset PIRB_HOST_COPY_PATH; fi
  for rep in 1 2; do
    python bench.py --steps 40 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$mode', 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'GB/s', round(d['e2e']['ms_per_step'],4), 'ms vs device', round(d['ms_per_step'],4))" | tee -a gpurun_out/e2e_ab.log
  done
  PIRB_BENCH_LOG_ROWS=20 python bench.py --steps 40 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$mode 2^20', 'e2e', round(d['e2e']['ms_per_step'],4), 'ms vs device', round(d['ms_per_step'],4))" | tee -a gpurun_out/e2e_ab.log
done
