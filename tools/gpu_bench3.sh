#!/bin/bash
# three bench lines of the tcgen05 FNO kernel set: c4 (256 members), c4 with 32 members (one rank's share at 8 GPUs), c2
mkdir -p gpurun_out
tag=${1:-x}
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c4_$tag.json 2> gpurun_out/r02_bench_$tag.err
python bench.py --members 32 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c4m32_$tag.json 2>> gpurun_out/r02_bench_$tag.err
python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c2_$tag.json 2>> gpurun_out/r02_bench_$tag.err
python bench.py --workload c1 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c1_$tag.json 2>> gpurun_out/r02_bench_$tag.err
tail -3 gpurun_out/r02_bench_$tag.err
for w in c4 c4m32 c2 c1; do python - gpurun_out/r02_bench_${w}_$tag.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split('/')[-1], round(d['value']), 'ms/rollout %.3f' % d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['kernel_ms_per_rollout'], 'step_frac %.3f' % d['step_roofline']['frac_of_hbm_peak'])
PY
done
