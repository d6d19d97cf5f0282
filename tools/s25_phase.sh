cp pir_b200/lib/variants/${V:-e}.so pir_b200/lib/libpir_b200.so
PIRB_PHASE_CASES="${CASES:-1:b4,1:b2}" bash tools/phase_debug.sh 2>&1 | grep -v "^\[cta" | tee gpurun_out/s25_phase_${V:-e}.log
