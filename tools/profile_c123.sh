set -x
python -m pytest tests -m gpu -q > gpurun_out/tests_s20.log 2>&1; echo "tests rc=$?"
for w in c1 c2 c3; do
python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
done
python bench.py --workload c3 --precision fp32 > gpurun_out/bench_c3_fp32.json 2>> gpurun_out/bench_c3.err; echo "rc=$?"
python bench.py --workload c3 --members 256 > gpurun_out/bench_c3_m256.json 2>> gpurun_out/bench_c3.err; echo "rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c3.csv python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_l_c3.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/tests_s20.log; cat gpurun_out/bench_c*.json
