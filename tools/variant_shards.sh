#!/bin/bash
# A/B of library builds in pir_b200/lib/variants/*.so on the shard probe (development aid): VARIANTS="pre_la cur"
mkdir -p gpurun_out
for v in ${VARIANTS:-pre_la cur}; do
  echo "=== variant $v" | tee -a gpurun_out/variant_shards.log
  VARIANT_SO=$PWD/pir_b200/lib/variants/$v.so python tools/shard_probe.py 2>&1 | tee -a gpurun_out/variant_shards.log
done
