#!/bin/bash
# on an 8-GPU box: host-memory ceiling probe at 1 / 2 / 4 / 8 ranks, strong-scaling bench lines at N = 1, 2, 4 (N = 8: tools/gpu_scale.sh)
mkdir -p gpurun_out
for n in 1 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2960$n tools/pcie_probe_ranks.py 2>/dev/null | grep ranks >> gpurun_out/r02_pcie_probe_ranks.txt
done
PROBE_MEMBERS=32 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29609 tools/pcie_probe_ranks.py 2>/dev/null | grep ranks >> gpurun_out/r02_pcie_probe_ranks.txt
cat gpurun_out/r02_pcie_probe_ranks.txt
python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_scale_strong_n1.json 2> gpurun_out/r02_scale_n1.err
for n in 2 4; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r02_scale_strong_n$n.json 2> gpurun_out/r02_scale_n$n.err
done
for n in 1 2 4; do python - gpurun_out/r02_scale_strong_n$n.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split('/')[-1], d['scaling'], 'n', d['n_gpus'], round(d['value']), 'ms %.3f' % d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['config']['members_per_gpu'], d['config']['members_total'])
PY
done
