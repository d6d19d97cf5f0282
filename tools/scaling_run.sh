#!/bin/bash
# gpurun --gpus 8 -- 'bash tools/scaling_run.sh': the multi-GPU tests, then bench.py at N = 1, 2, 4, 8 the way the driver
# launches it (torchrun for N > 1), both arms at N = 1.  V names the outputs.
V=${V:-r2}
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_golden_fullsize.py::test_current_device_is_left_alone -m gpu -q 2>&1 | tail -3
for N in ${NS:-1 2 4 8}; do
  if [ $N = 1 ]; then CMD="python bench.py --gpus 1"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N"; fi
  ( time timeout 900 $CMD --steps ${STEPS:-20} --warmup 5 ) > gpurun_out/${V}_scale_n$N.json 2> gpurun_out/${V}_scale_n$N.err; echo "N=$N rc=$?"; tail -4 gpurun_out/${V}_scale_n$N.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${V}_scale_n$N.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('N=$N value',round(d['value']),'ms',round(d['ms_per_step'],4),'e2e',round(d['e2e']['value']),round(d['e2e']['ms_per_step'],4),'one-at-a-time',round(d['e2e_one_at_a_time']['value']),round(d['e2e_one_at_a_time']['ms_per_step'],4),'sp',d.get('e2e_single_process',{}).get('ms_per_step'),'iso',round(r['kernel_ms_isolated'],4),'aes',round(r['lookup']['aes_only_ms'],4),'checked',d['parity_checked'])
    for o in d.get('other_configs',[]): print('  ',o['config'], round(o['ms_per_step'],4), round(o['value'],1), o['parity_checked'])
except Exception as e: print('parse failed',e)
PY
done
