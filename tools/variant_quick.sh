# one library variant: parity + multi-GPU tests, shard sweep, 2^20 x 256 B (development aid)
mkdir -p gpurun_out
cp pir_b200/lib/variants/$V.so pir_b200/lib/libpir_b200.so
echo "=== variant $V" | tee -a gpurun_out/variant_quick.log
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -3 | tee -a gpurun_out/variant_quick.log
SHAPES="auto" timeout 600 python tools/split_sweep.py 2>&1 | tee -a gpurun_out/variant_quick.log
LOG_ROWS=20 SHAPES="auto" timeout 600 python tools/split_sweep.py 2>&1 | grep "1/1" | tee -a gpurun_out/variant_quick.log
