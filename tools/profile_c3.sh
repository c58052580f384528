# Evidence for the AFNO (c3) kernel set: bench lines at 8 and 256 members, launch list of one rollout,
# full capture of one model evaluation (4 kernels x 4 blocks) at 256 members.
set -x
python bench.py --workload c3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "rc=$?"
python bench.py --workload c3 --members 256 > gpurun_out/bench_c3_m256.json 2>> gpurun_out/bench_c3.err; echo "rc=$?"
python bench.py --workload c3 --precision fp32 --no-cpu-baseline > gpurun_out/bench_c3_fp32.json 2>> gpurun_out/bench_c3.err; echo "rc=$?"
python bench.py --workload c3 --members 256 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c3_m256.csv python bench.py --workload c3 --members 256 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l_c3.log 2>&1; echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:afno -s 16 -c 16 -o gpurun_out/afno_c3_m256 -f python bench.py --workload c3 --members 256 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f_c3.log 2>&1; echo "ncu2 rc=$?"
