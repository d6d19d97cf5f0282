#!/bin/bash
# what the driver runs at round end, in one call: GPU tests, smoke(), bench (both arms)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | cut -c1-160
timeout 600 python bench.py 2>/dev/null | tee gpurun_out/bench_final_default.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('bench', round(d['value']), 'GB/s', round(d['ms_per_step'],4), 'ms e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],3), 'launches', d['gpu_launches'])"
