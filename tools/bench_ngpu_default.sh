#!/bin/bash
# bench.py on N GPUs exactly as the driver launches it (default legs): tools/bench_ngpu_default.sh N
O=gpurun_out; mkdir -p $O
N=${1:-2}
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 12 --warmup 3 > $O/bench_default_$N.json 2> $O/bench_default_$N.err; echo "bench$N rc=$?"; tail -n 3 $O/bench_default_$N.err | cut -c1-300
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_default_$N.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','gpu_launches','n_gpus')}); print('e2e', {k: d['e2e'][k] for k in ('value','ms_per_step','h2d_bytes_per_step','d2h_bytes_per_step','coherence')}, 'strict', d['e2e'].get('strict',{}).get('value'))
print('parity', d.get('multi_rank_parity',{}).get('ok'), 'yee', d['yee']['field_step']['cell_updates_per_s'], 'clocks', d['clocks'])
PY
