mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 | tee gpurun_out/s25_final.log
python bench.py --steps 40 --warmup 5 2>/dev/null | tee gpurun_out/s25_bench40.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('headline', round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), 'frac', d['roofline']['frac'])" | tee -a gpurun_out/s25_final.log
bash tools/batch256_run.sh 2>&1 | tee -a gpurun_out/s25_final.log
PIRB_BENCH_LOG_ROWS=22 PIRB_BENCH_SLOT_BYTES=1024 PIRB_BENCH_QUERIES=64 python bench.py --steps 5 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('c3 batch64', 'GB/s', round(d['value']), 'ms/batch', round(d['ms_per_step'],3), 'q/s', round(d['queries_per_s'],1))" | tee -a gpurun_out/s25_final.log
PIRB_BENCH_LOG_ROWS=20 python bench.py --steps 40 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('2^20x256', 'GB/s', round(d['value']), 'ms', round(d['ms_per_step'],4))" | tee -a gpurun_out/s25_final.log
