# A/B of the batch scan kernel (two selected rows per step against one), configs[3] shape and 256 x 2^20 x 1 KiB
mkdir -p gpurun_out
LOG=gpurun_out/r2_scan_ab.txt
: > $LOG
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden_fullsize.py -x -q -m gpu -k "batch or 64" 2>&1 | tail -3 | tee -a $LOG
for rep in 1 2; do
  for v in scan_old_dev scan_new_dev; do
    echo "=== $v (64 x 2^22 x 1 KiB)" | tee -a $LOG
    DEV_SO=pir_b200/lib/variants/$v.so timeout 300 python tools/batch_phases.py 2>&1 | grep bits | tee -a $LOG
  done
done
for v in scan_old_dev scan_new_dev; do
  echo "=== $v (64 x 2^20 x 4 KiB)" | tee -a $LOG
  DEV_SO=pir_b200/lib/variants/$v.so LOG_ROWS=20 SLOT=4096 timeout 300 python tools/batch_phases.py 2>&1 | grep bits | tee -a $LOG
  echo "=== $v (64 x 2^22 x 528 B)" | tee -a $LOG
  DEV_SO=pir_b200/lib/variants/$v.so LOG_ROWS=22 SLOT=528 timeout 300 python tools/batch_phases.py 2>&1 | grep bits | tee -a $LOG
done
