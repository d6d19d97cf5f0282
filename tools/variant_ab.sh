# A/B of library variants built into pir_b200/lib/variants (development aid)
mkdir -p gpurun_out
cp pir_b200/lib/libpir_b200.so /tmp/orig.so
for v in ${VARIANTS:-v0 a b}; do
  cp pir_b200/lib/variants/$v.so pir_b200/lib/libpir_b200.so
  echo "=== variant $v" | tee -a gpurun_out/variant_ab.log
  if [ "$v" != "v0" ]; then
    timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -3 | tee -a gpurun_out/variant_ab.log
  fi
  SHAPES="${SHAPES:-auto b4 b3 b2}" timeout 600 python tools/split_sweep.py 2>&1 | tee -a gpurun_out/variant_ab.log
done
cp /tmp/orig.so pir_b200/lib/libpir_b200.so
