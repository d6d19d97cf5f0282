#!/bin/bash
# 2 GPUs: the multi-GPU tests (CUDA IPC between two processes), bench at N=1 and N=2 (torchrun), reference arm at N=2
mkdir -p gpurun_out
nvidia-smi -L | head -3
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_golden_fullsize.py::test_current_device_is_left_alone -m gpu -q 2>&1 | tail -3
for N in 1 2; do
  if [ $N = 1 ]; then CMD="python bench.py --gpus 1"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N"; fi
  ( time timeout 900 $CMD --steps 20 --warmup 5 ) > gpurun_out/r2_scale_n$N.json 2> gpurun_out/r2_scale_n$N.err; echo "N=$N rc=$?"; tail -4 gpurun_out/r2_scale_n$N.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2_scale_n$N.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('N=$N value',round(d['value']),'ms',round(d['ms_per_step'],4),'e2e',round(d['e2e']['value']),round(d['e2e']['ms_per_step'],4),'sp',d.get('e2e_single_process',{}).get('ms_per_step'),'iso',round(r['kernel_ms_isolated'],4),'aes',round(r['lookup']['aes_only_ms'],4),'checked',d['parity_checked'])
    for o in d.get('other_configs',[]): print('  ',o['config'], round(o['ms_per_step'],4), round(o['value'],1), o['parity_checked'])
except Exception as e: print('parse failed',e)
PY
done
