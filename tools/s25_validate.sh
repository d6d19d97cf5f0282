set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s25_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/s25_pytest.log
tail -5 gpurun_out/s25_pytest.log
timeout 300 python bench.py > gpurun_out/s25_bench.json 2> gpurun_out/s25_bench.err; echo "bench exit $?"
cat gpurun_out/s25_bench.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/s25_bench_ref.json 2> gpurun_out/s25_bench_ref.err; echo "ref exit $?"
cat gpurun_out/s25_bench_ref.json
which go; go version
