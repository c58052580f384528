# End-of-round evidence on one GPU: GPU test log, the bench line (c4) and the reference arm, parity-config lines.
set -x
python -m pytest tests -m gpu -q > gpurun_out/tests_v4.log 2>&1; echo "tests rc=$?"
python bench.py > gpurun_out/bench_v4.json 2> gpurun_out/bench_v4.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref_v4.json 2>> gpurun_out/bench_v4.err; echo "ref rc=$?"
for w in c1 c2; do python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_${w}_v4.json 2>> gpurun_out/bench_v4.err; echo "rc=$?"; done
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_v4.log 2>&1; echo "smoke rc=$?"
tail -3 gpurun_out/tests_v4.log; cut -c1-250 gpurun_out/bench_v4.json
