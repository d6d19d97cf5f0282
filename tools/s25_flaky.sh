mkdir -p gpurun_out
for v in e3o1 e3o4; do
  cp pir_b200/lib/variants/$v.so pir_b200/lib/libpir_b200.so
  echo "=== $v" | tee -a gpurun_out/s25_flaky.log
  for i in 1 2 3 4 5 6; do
    timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "two_phase" 2>&1 | tail -2 | tee -a gpurun_out/s25_flaky.log
  done
done
