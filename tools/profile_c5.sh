set -x
python bench.py --workload c5 --precision bf16 --steps 3 --warmup 3 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench rc=$?"
cat gpurun_out/bench_c5.json
python bench.py --workload c5 --precision fp32 --members 1 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_fp32.json 2>> gpurun_out/bench_c5.err; echo "bench fp32 rc=$?"
python bench.py --workload c5 --precision bf16 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c5.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tcw_|mix_dense|wide_gather|metrics" -s 69 -c 185 --csv --log-file gpurun_out/launches_c5.csv python bench.py --workload c5 --precision bf16 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l_c5.log 2>&1; echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"tcw_|mix_dense|wide_gather" -s 23 -c 23 -o gpurun_out/prof_c5 -f python bench.py --workload c5 --precision bf16 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f_c5.log 2>&1; echo "ncu2 rc=$?"
