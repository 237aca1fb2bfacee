#!/bin/bash
# Round 2, GPU session g (4 GPUs): multi-GPU tests on real devices (NCCL + put transports), N-rank decks with one GPU per rank, bench at N=4
O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8
timeout 2400 python -m pytest tests/test_multi_rank.py -q -m gpu > $O/r2g_multirank.log 2>&1; echo "multirank rc=$?"; tail -n 12 $O/r2g_multirank.log
timeout 2400 python -m pytest tests/test_gpu_dropin.py -q -m gpu -k "n_ranks or pcomm" > $O/r2g_dropin.log 2>&1; echo "dropin rc=$?"; tail -n 8 $O/r2g_dropin.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 4 --steps 8 --warmup 3 > $O/r2g_bench4.json 2> $O/r2g_bench4.err; echo "bench4 rc=$?"; tail -n 3 $O/r2g_bench4.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2g_bench4.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','gpu_launches','n_gpus')}); print('roofline', d['roofline']['frac'], d['roofline']['ms_per_launch_by_species'], 'sort', d['sort_p']['ms_per_call'])
print('parity', d.get('multi_rank_parity')); print('yee', d.get('yee',{}).get('field_step'))
print('e2e', {k: d['e2e'][k] for k in d['e2e'] if k not in ('workload','energies_after')})
PY
ls -la $O | grep r2g
