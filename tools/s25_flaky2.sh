mkdir -p gpurun_out; : > gpurun_out/s25_flaky2.log
for i in $(seq 1 ${N:-20}); do
  timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -x 2>&1 | tail -40 > /tmp/one.log
  tail -1 /tmp/one.log | tee -a gpurun_out/s25_flaky2.log
  if grep -q failed /tmp/one.log; then cat /tmp/one.log >> gpurun_out/s25_flaky2.log; fi
done
