#!/bin/bash
# round 2, first lease: the GPU test-suite (new steady-state parity cases), the no-alloc test alone, a bench line of the
# code as it stands, and compute-sanitizer over the three hand-written pipeline families
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s 2>&1 | tail -150 > gpurun_out/r02_tests_gpu_a.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -s > gpurun_out/r02_tests_zz_alone.log 2>&1
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_start.json 2> gpurun_out/r02_bench_start.err
S=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck synccheck racecheck; do
  for c in c4 afno wide; do
    timeout 420 $S --tool $tool --error-exitcode 7 --print-limit 20 python tools/sanitize.py $c > gpurun_out/r02_sanitize_${tool}_${c}.log 2>&1
    echo "rc=$?" >> gpurun_out/r02_sanitize_${tool}_${c}.log
  done
done
tail -5 gpurun_out/r02_tests_gpu_a.log; tail -3 gpurun_out/r02_tests_zz_alone.log; for f in gpurun_out/r02_sanitize_*.log; do echo $f; tail -4 $f; done
