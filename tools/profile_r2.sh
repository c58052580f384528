#!/bin/bash
# Round-2 evidence on one GPU: GPU test log, bench lines (c4 with cpu / eager / parity legs, reference arm, c1-c3, c4 at 32 members),
# smoke, then under ncu the launch list of one rollout and --set full of one evaluation (c4).   usage: profile_r2.sh TAG
tag=${1:-a}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s > gpurun_out/r02_tests_gpu_$tag.log 2>&1; echo "tests rc=$?"; tail -1 gpurun_out/r02_tests_gpu_$tag.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_c4_$tag.json 2> gpurun_out/r02_bench_$tag.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/r02_bench_reference_arm_$tag.json 2>> gpurun_out/r02_bench_$tag.err; echo "ref rc=$?"
for w in c1 c2 c3; do python bench.py --workload $w > gpurun_out/r02_bench_${w}_$tag.json 2>> gpurun_out/r02_bench_$tag.err; echo "$w rc=$?"; done
python bench.py --members 32 --no-cpu-baseline > gpurun_out/r02_bench_c4m32_$tag.json 2>> gpurun_out/r02_bench_$tag.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_smoke_$tag.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain_$tag.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 620 -c 201 --csv --log-file gpurun_out/r02_ncu_launches_c4_$tag.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l_$tag.log 2>&1; echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"fno_layer|fno_lift|fno_mid|fno_proj" -s 20 -c 10 -o gpurun_out/r02_prof_c4_$tag -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f_$tag.log 2>&1; echo "ncu2 rc=$?"
python tools/ncu_summary_csv.py gpurun_out/r02_prof_c4_$tag.ncu-rep gpurun_out/r02_ncu_full_c4_$tag.csv > /dev/null 2>&1; echo "summary rc=$?"
tail -3 gpurun_out/r02_tests_gpu_$tag.log; cut -c1-200 gpurun_out/r02_bench_c4_$tag.json
# AFNO (config c3): bench lines at 8 (above) and 256 members, A/B against the fp32 mid kernel, --set full of one block's four kernels
python bench.py --workload c3 --members 256 --no-cpu-baseline > gpurun_out/r02_bench_c3m256_$tag.json 2>> gpurun_out/r02_bench_$tag.err; echo "c3m256 rc=$?"
DLWP_MID_FMA=1 python bench.py --workload c3 --members 256 --no-cpu-baseline --no-eager > gpurun_out/r02_bench_c3m256_midfma_$tag.json 2>> gpurun_out/r02_bench_$tag.err
ncu --set full --clock-control none --import-source on -k regex:afno -s 20 -c 4 -o gpurun_out/r02_prof_c3_$tag -f python bench.py --workload c3 --members 256 --steps 1 --warmup 3 --no-cpu-baseline --no-eager > gpurun_out/ncu_f_c3_$tag.log 2>&1; echo "ncu3 rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:afno -c 100 --csv --log-file gpurun_out/r02_ncu_launches_c3_$tag.csv python bench.py --workload c3 --members 256 --steps 1 --warmup 3 --no-cpu-baseline --no-eager > /dev/null 2>&1; echo "ncu4 rc=$?"
(python tools/ncu_pipe.py gpurun_out/r02_prof_c4_$tag.ncu-rep; python tools/ncu_pipe.py gpurun_out/r02_prof_c3_$tag.ncu-rep) > gpurun_out/r02_data_pipe_$tag.txt 2>&1
tools/ubench/umma > gpurun_out/r02_ubench_umma_$tag.txt 2>&1; tools/ubench/umma_ts >> gpurun_out/r02_ubench_umma_$tag.txt 2>&1
tools/ubench/gelu_rate > gpurun_out/r02_ubench_gelu_$tag.txt 2>&1
# config 5 on the bf16 kernel set, and the same with the CUDA-core few-sample mode mixing
python bench.py --workload c5 --precision bf16 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c5_bf16_$tag.json 2>> gpurun_out/r02_bench_$tag.err; echo "c5 rc=$?"
DLWP_MIX_FMA=1 python bench.py --workload c5 --precision bf16 --steps 3 --warmup 3 --no-cpu-baseline --no-eager > gpurun_out/r02_bench_c5_bf16_mixfma_$tag.json 2>> gpurun_out/r02_bench_$tag.err
DLWP_LIFT4=0 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-eager > gpurun_out/r02_bench_c4_lift2_$tag.json 2>> gpurun_out/r02_bench_$tag.err
DLWP_L2_HINTS=0 DLWP_LAYER_REV=0 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-eager > gpurun_out/r02_bench_c4_nol2_$tag.json 2>> gpurun_out/r02_bench_$tag.err
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-eager > gpurun_out/r02_bench_c4_again_$tag.json 2>> gpurun_out/r02_bench_$tag.err
cat gpurun_out/r02_data_pipe_$tag.txt
