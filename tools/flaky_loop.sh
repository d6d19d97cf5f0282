# loop the batch tests inside pytest and keep the assertion message of any failure (development aid)
mkdir -p gpurun_out; : > gpurun_out/flaky_loop.log
for i in $(seq 1 ${N:-40}); do
  timeout 120 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "repeatable or two_phase" 2>&1 > /tmp/one.log
  tail -1 /tmp/one.log >> gpurun_out/flaky_loop.log
  if grep -q failed /tmp/one.log; then grep -E "assert|Error|rep|mode|FAILED" /tmp/one.log | head -20 >> gpurun_out/flaky_loop.log; fi
done
sort gpurun_out/flaky_loop.log | uniq -c | sort -rn | head -12
