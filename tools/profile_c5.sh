set -x
python -m pytest tests -m gpu -q -x > gpurun_out/tests_c5.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/tests_c5.log
python bench.py --workload c5 --precision bf16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench rc=$?"
cat gpurun_out/bench_c5.json
ncu --set full --clock-control none --import-source on -k regex:"tcw_kernel|tdft|mix_dense|wide_" -s 27 -c 27 -o gpurun_out/prof_c5 -f python bench.py --workload c5 --precision bf16 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c5.log 2>&1; echo "ncu rc=$?"
