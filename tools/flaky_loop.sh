#!/bin/bash
# Stress of the round-1 intermittent defect (wrong database on padded slot sizes after an
# unordered cudaMemset): N fresh PROCESSES, P at a time, each creating fresh handles of 100-byte slots, filling,
# and answering a batch through both batch paths against the oracle.  Usage: N=300 P=6 tools/flaky_loop.sh
mkdir -p gpurun_out
N=${N:-300}; P=${P:-6}
LOG=gpurun_out/flaky_loop.log; : > $LOG
one() {
  timeout 300 python tools/flaky_one.py "$1" >> $LOG 2>&1 || echo "start $1: FAILED rc=$?" >> $LOG
}
export -f one; export LOG
seq 1 $N | xargs -P $P -I{} bash -c 'one {}'
echo "process starts: $(grep -c '^start' $LOG)  ok: $(grep -c ' ok$' $LOG)  failed: $(grep -c FAILED $LOG)  mismatches: $(grep -c MISMATCH $LOG)"
