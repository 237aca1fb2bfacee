O=gpurun_out; mkdir -p $O
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 "$@" 2>$O/c6_err.log | tail -n 1; }
run --cells 128x128x128 --steps 16 --warmup 4 --no-cpu --no-e2e --no-yee --breakdown 16 > $O/c6_2gpu.json
python bench.py --cells 128x128x128 --steps 16 --warmup 4 --no-cpu --no-e2e --no-yee --breakdown 16 > $O/c6_1gpu.json
python - <<'PY'
import json
for f in ("gpurun_out/c6_1gpu.json","gpurun_out/c6_2gpu.json"):
    d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, "ms/step %.3f"%d["ms_per_step"], d["breakdown_ms"])
PY
