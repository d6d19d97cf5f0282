mkdir -p gpurun_out
cp pir_b200/lib/variants/${V:-w}.so pir_b200/lib/libpir_b200.so
for wt in ${WTS:-5 3}; do
  echo "=== PIRB_WT=$wt pytest" | tee -a gpurun_out/s25_wt.log
  PIRB_WT=$wt timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -15 | tee -a gpurun_out/s25_wt.log
done
SHAPES="${SHAPES:-auto w4 w3 w2}" timeout 600 python tools/split_sweep.py 2>&1 | tee -a gpurun_out/s25_wt.log
