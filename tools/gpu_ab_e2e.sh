# as gpu_ab.sh, printing the end-to-end (host buffers) figure too
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
print(sys.argv[1], "device", round(d["value"]), round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"], 4))
PY
  done
done
cp /tmp/lib_keep.so $L
