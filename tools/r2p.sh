#!/bin/bash
# key patching in boundary_p: parity (1 GPU lazy tests, N ranks on one GPU resident) + 2-GPU bench
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "lazy or unprocessed" > $O/r2p_lazy.log 2>&1; echo "lazy rc=$?"; tail -n 3 $O/r2p_lazy.log
timeout 1500 python -m pytest tests/test_multi_rank.py -x -q -m gpu -k "resident or two_gpus_match" > $O/r2p_mr.log 2>&1; echo "mr rc=$?"; tail -n 30 $O/r2p_mr.log | cut -c1-400
N=2
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 12 --warmup 3 --no-e2e --breakdown 6 > $O/r2p_bench$N.json 2> $O/r2p_bench$N.err; echo "bench$N rc=$?"; tail -n 2 $O/r2p_bench$N.err | cut -c1-300
python - <<PY
import json
d=json.loads(open('gpurun_out/r2p_bench$N.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','gpu_launches','n_gpus')}); print('roofline', d['roofline']['frac'], d['roofline']['ms_per_launch_by_species'], 'sort', d['sort_p'])
print('parity', d.get('multi_rank_parity')); print('breakdown', d.get('breakdown_ms'))
PY
