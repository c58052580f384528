# AFNO (c3) check: parity tests, bench at 8 and 256 members, per-kernel profile lines
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
for m in 0 256; do
timeout 200 python bench.py --workload c3 --members $m --no-cpu-baseline > gpurun_out/c3_m$m.json 2> gpurun_out/c3_m$m.err
python - <<PY
import json
j=json.loads(open('gpurun_out/c3_m$m.json').read().strip().split('\n')[-1])
print($m, round(j['value']), j['ms_per_step'], j['kernel_ms_per_rollout'])
PY
done
