mkdir -p gpurun_out
cp pir_b200/lib/variants/${V:-q}.so pir_b200/lib/libpir_b200.so
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/s25_q.log
SHAPES="auto" timeout 600 python tools/split_sweep.py 2>&1 | tee -a gpurun_out/s25_q.log
LOG_ROWS=20 SHAPES="auto" timeout 600 python tools/split_sweep.py 2>&1 | grep "1/1" | tee -a gpurun_out/s25_q.log
PIRB_PHASE_CASES="1:auto,8:auto" bash tools/phase_debug.sh 2>&1 | grep -E "shard|ctas" | tee -a gpurun_out/s25_q.log
