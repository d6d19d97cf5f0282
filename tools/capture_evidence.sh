set -x
python bench.py --steps 40 --warmup 5 > gpurun_out/bench_v9.json 2> gpurun_out/bench_v9.err; echo rc=$?
python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_v9_ref.json 2> gpurun_out/bench_v9_ref.err; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v9.csv python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_v9.log 2>&1; echo rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:answer_kernel -c 2 -o gpurun_out/answer_v9 -f python tools/parity_probe.py > gpurun_out/ncu_full_v9.log 2>&1; echo rc=$?
ls -la gpurun_out/ | tail -8
