#!/bin/bash
# A/B of the texture-path row fetch (-DPIRB_TEX_ROWS build in pir_b200/lib/variants/tex.so) against the shipped build
mkdir -p gpurun_out
for v in cur tex cur tex; do
  echo "=== variant $v"; VARIANT_SO=$PWD/pir_b200/lib/variants/$v.so python tools/parity_probe.py 2>&1 | grep -v "^\[tex\]"
done | tee gpurun_out/r2_tex_ab.txt
for v in cur tex; do
  timeout 600 ncu --set full --clock-control none -k regex:answer_kernel --launch-skip 12 -c 1 -o gpurun_out/answer_${v}_odd -f env VARIANT_SO=$PWD/pir_b200/lib/variants/$v.so python tools/parity_probe.py > gpurun_out/ncu_${v}.log 2>&1; echo "ncu $v rc=$?"
done
