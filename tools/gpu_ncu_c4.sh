# ncu --set full with source import for the c4 MLP kernels (one launch each)
set -x
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c4.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${1:-fno_proj|fno_lift}" -s ${2:-2} -c ${3:-2} -o gpurun_out/prof_c4x -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c4x.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/prof_c4x.ncu-rep
