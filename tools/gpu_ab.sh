# A/B of library variants on the same box: dlwp-benchmark_b200/_variants/lib{A,B,...}.so are swapped in turn
# usage: gpu_ab.sh "<bench args>" A B [C ...]
args="$1"; shift
L=dlwp-benchmark_b200/libdlwp_b200.so
cp $L /tmp/lib_keep.so
for rep in 1 2 3; do
  for v in "$@"; do
    cp dlwp-benchmark_b200/_variants/lib$v.so $L
    timeout 600 python bench.py $args --no-cpu-baseline > gpurun_out/ab_$v.json 2> gpurun_out/ab.err || { echo "$v failed"; tail -5 gpurun_out/ab.err; }
    python - "$v" <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/ab_{sys.argv[1]}.json").read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"]), round(d["ms_per_step"], 4), d.get("kernel_ms_per_rollout"))
PY
  done
done
cp /tmp/lib_keep.so $L
