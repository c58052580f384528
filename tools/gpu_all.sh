#!/bin/bash
# full GPU test suite, then the bench lines c4 (256 and 32 members), c2, c1, c3 (8 and 256 members).  usage: tools/gpu_all.sh TAG
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/tests_$TAG.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/tests_$TAG.log
run() { name=$1; shift; timeout 300 python bench.py "$@" --no-cpu-baseline --no-eager > gpurun_out/bench_${name}_$TAG.json 2> gpurun_out/bench_${name}_$TAG.err || tail -3 gpurun_out/bench_${name}_$TAG.err
  python - gpurun_out/bench_${name}_$TAG.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], round(d['value']), 'ms/rollout %.3f' % d['ms_per_step'], 'e2e', round(d['e2e']['value']), d.get('kernel_ms_per_rollout'), 'parity', d.get('parity', {}).get('max_rel_increment') if isinstance(d.get('parity'), dict) else None)
except Exception as e:
    print(sys.argv[1], 'failed', e)
PY
}
run c4 --steps 10 --warmup 3
run c4m32 --members 32 --steps 10 --warmup 3
run c2 --workload c2 --steps 10 --warmup 3
run c1 --workload c1 --steps 5 --warmup 3
run c3 --workload c3
run c3m256 --workload c3 --members 256
