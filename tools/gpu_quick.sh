# Quick A/B on one GPU: the touched tests, then the c4 bench line and c3 at 256 members.
set -x
timeout 900 python -m pytest tests -m gpu -q -x -k "metrics or tensor_core or afno" > gpurun_out/quick_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/quick_tests.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/quick_c4.json 2> gpurun_out/quick.err; echo "c4 rc=$?"
timeout 600 python bench.py --workload c3 --members 256 --no-cpu-baseline > gpurun_out/quick_c3_m256.json 2>> gpurun_out/quick.err; echo "c3 rc=$?"
python - <<'PY'
import json
for f in ("quick_c4", "quick_c3_m256"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["value"]), d["ms_per_step"], d.get("kernel_ms_per_rollout"))
    except Exception as e:
        print(f, "failed", e)
PY
