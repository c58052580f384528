# c4 bench line only (no CPU leg), kernel times per rollout; optional first argument: pytest -k filter to run before
set -x
if [ -n "$1" ]; then timeout 900 python -m pytest tests -m gpu -q -x -k "$1" 2>&1 | tail -3; fi
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/quick_c4.json 2> gpurun_out/quick.err; echo "c4 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/quick_c4.json").read().strip().splitlines()[-1])
print("c4", round(d["value"]), d["ms_per_step"], "e2e", round(d["e2e"]["value"]), d.get("kernel_ms_per_rollout"))
PY
