#!/bin/bash
# AFNO (config c3) on the GPU box: parity tests, then the c3 bench at 8 and 256 members with the tcgen05 mid kernel and
# with the fp32 FMA mid kernel (DLWP_MID_FMA=1) for an A/B.   usage: tools/gpu_afno.sh TAG
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -q -m gpu -x -k "afno" > gpurun_out/afno_tests_$TAG.log 2>&1
echo "tests rc=$?" >> gpurun_out/afno_tests_$TAG.log
tail -5 gpurun_out/afno_tests_$TAG.log
for m in 8 256; do
  timeout 300 python bench.py --workload c3 --members $m --no-cpu-baseline --no-eager > gpurun_out/afno_c3_m${m}_$TAG.json 2> gpurun_out/afno_c3_m${m}_$TAG.err
  DLWP_MID_FMA=1 timeout 300 python bench.py --workload c3 --members $m --no-cpu-baseline --no-eager > gpurun_out/afno_c3_m${m}_fma_$TAG.json 2>> gpurun_out/afno_c3_m${m}_$TAG.err
  tail -2 gpurun_out/afno_c3_m${m}_$TAG.err
done
for n in afno_c3_m8_$TAG afno_c3_m8_fma_$TAG afno_c3_m256_$TAG afno_c3_m256_fma_$TAG; do python - gpurun_out/$n.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], round(d['value']), 'ms/rollout %.3f' % d['ms_per_step'], d.get('kernel_ms_per_rollout'), d.get('parity'))
except Exception as e:
    print(sys.argv[1], 'failed', e)
PY
done
