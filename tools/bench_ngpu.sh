#!/bin/bash
# bench.py on N GPUs of one node (torchrun), reduced legs + per-entry-point breakdown: tools/bench_ngpu.sh N
O=gpurun_out; mkdir -p $O
N=${1:-8}
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 12 --warmup 3 --no-e2e --breakdown 6 > $O/bench_ngpu_$N.json 2> $O/bench_ngpu_$N.err; echo "bench$N rc=$?"; tail -n 2 $O/bench_ngpu_$N.err | cut -c1-300
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_ngpu_$N.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','gpu_launches','n_gpus')}); print('roofline', d['roofline']['frac'], d['roofline']['ms_per_launch_by_species'], 'sort', d['sort_p']['ms_per_call'])
print('parity', d.get('multi_rank_parity')); print('yee', d.get('yee',{}).get('field_step')); print('breakdown', d.get('breakdown_ms'))
PY
