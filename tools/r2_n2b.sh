#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/variant_shards.log
CUDA_VISIBLE_DEVICES=0 VARIANTS="cur2 widebfs" bash tools/variant_shards.sh
N=2
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 ) > gpurun_out/r2_scale_n$N.json 2> gpurun_out/r2_scale_n$N.err; echo "N=$N rc=$?"; tail -4 gpurun_out/r2_scale_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_scale_n$N.json').read().strip().splitlines()[-1]); r=d['roofline']
print('N=$N value',round(d['value']),'ms',round(d['ms_per_step'],4),'e2e',round(d['e2e']['value']),round(d['e2e']['ms_per_step'],4),'sp',d.get('e2e_single_process',{}).get('ms_per_step'),'iso',round(r['kernel_ms_isolated'],4),'aes',round(r['lookup']['aes_only_ms'],4),'checked',d['parity_checked'])
for o in d.get('other_configs',[]): print('  ',o['config'], round(o['ms_per_step'],4), round(o['value'],1), o['parity_checked'])
PY
