#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py tests/test_gpu_golden_fullsize.py -m gpu -q 2>&1 | tail -3
python tools/parity_probe.py
python tools/shard_probe.py 2>&1 | tee gpurun_out/r2_shard_probe_v17.txt
python bench.py > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2f_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2f_bench.json')); r=d['roofline']
print('value',round(d['value']),'ms',round(d['ms_per_step'],4),'e2e',round(d['e2e']['value']),round(d['e2e']['ms_per_step'],4),'iso',round(r['kernel_ms_isolated'],4),'aes',round(r['lookup']['aes_only_ms'],4))
PY
