#!/bin/bash
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -q ) > gpurun_out/r2c_gpu_tests.log 2>&1; tail -8 gpurun_out/r2c_gpu_tests.log
python tools/shard_probe.py 2>&1 | tee gpurun_out/r2_shard_probe.txt
python bench.py > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo bench rc=$?; tail -3 gpurun_out/r2c_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2c_bench.json')); r=d['roofline']
print('value',round(d['value']),'ms',round(d['ms_per_step'],4),'e2e',round(d['e2e']['value']),'iso',round(r['kernel_ms_isolated'],4),'aes',round(r['lookup']['aes_only_ms'],4))
for o in d['other_configs']: print(o['config'], round(o['ms_per_step'],4), round(o['value'],1), o['parity_checked'])
PY
