set -x
python -m pytest tests -m gpu -q > gpurun_out/tests_v3.log 2>&1; echo "tests rc=$?"
python bench.py > gpurun_out/bench_v3.json 2> gpurun_out/bench_v3.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref_v3.json 2>> gpurun_out/bench_v3.err; echo "ref rc=$?"
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain_v3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 603 -c 201 --csv --log-file gpurun_out/launches_v3.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l_v3.log 2>&1; echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"fno_layer|fno_lift|fno_mid|fno_proj" -s 20 -c 10 -o gpurun_out/prof_v3 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f_v3.log 2>&1; echo "ncu2 rc=$?"
tail -3 gpurun_out/tests_v3.log; cat gpurun_out/bench_v3.json
