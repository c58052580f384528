#!/bin/bash
# strong-scaling line of workload c4 on N GPUs of one box (256 members sharded), and the weak-scaling line beside it
N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_scale_strong_n$N.json 2> gpurun_out/r02_scale_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --scaling weak --members 256 > gpurun_out/r02_scale_weak_n$N.json 2>> gpurun_out/r02_scale_n$N.err
tail -3 gpurun_out/r02_scale_n$N.err
for k in strong weak; do python - gpurun_out/r02_scale_${k}_n$N.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split('/')[-1], d['scaling'], 'n', d['n_gpus'], round(d['value']), 'ms %.3f' % d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['config']['members_per_gpu'], d['config']['members_total'], d['clocks'])
PY
done
