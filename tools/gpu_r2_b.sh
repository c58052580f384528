#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/r02_tests_gpu_b.log
tail -3 gpurun_out/r02_tests_gpu_b.log
python tools/graph_probe.py > gpurun_out/r02_graph_probe.txt 2>&1; cat gpurun_out/r02_graph_probe.txt
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_c4_b.json 2> gpurun_out/r02_bench_c4_b.err; tail -c 1500 gpurun_out/r02_bench_c4_b.json; tail -5 gpurun_out/r02_bench_c4_b.err
python bench.py --workload c2 --steps 10 --warmup 3 > gpurun_out/r02_bench_c2_b.json 2> gpurun_out/r02_bench_c2_b.err; tail -c 1200 gpurun_out/r02_bench_c2_b.json
python bench.py --members 32 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c4_m32.json 2>&1; head -c 300 gpurun_out/r02_bench_c4_m32.json
