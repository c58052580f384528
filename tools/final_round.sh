# End-of-round evidence on one GPU (v5): GPU test log, the bench line (c4), the reference arm, the parity-config
# lines, then under ncu the launch list of one rollout and --set full of one evaluation.
set -x
python -m pytest tests -m gpu -q > gpurun_out/tests_v5.log 2>&1; echo "tests rc=$?"
python bench.py > gpurun_out/bench_v5.json 2> gpurun_out/bench_v5.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref_v5.json 2>> gpurun_out/bench_v5.err; echo "ref rc=$?"
for w in c1 c2 c3; do python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_${w}_v5.json 2>> gpurun_out/bench_v5.err; echo "rc=$?"; done
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_v5.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain_v5.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 603 -c 201 --csv --log-file gpurun_out/launches_v5.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l_v5.log 2>&1; echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"fno_layer|fno_lift|fno_mid|fno_proj" -s 20 -c 10 -o gpurun_out/prof_v5 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f_v5.log 2>&1; echo "ncu2 rc=$?"
tail -3 gpurun_out/tests_v5.log; cut -c1-250 gpurun_out/bench_v5.json
