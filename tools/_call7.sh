O=gpurun_out; mkdir -p $O
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 "$@" 2>$O/c7_err.log | tail -n 1; }
run --steps 8 --warmup 3 --no-cpu --no-e2e --no-yee --breakdown 8 > $O/c7_2gpu_full.json
python - <<'PY'
import json
for f in ("gpurun_out/c7_2gpu_full.json",):
    d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, "Gpush/s %.2f ms/step %.3f"%(d["value"]/1e9,d["ms_per_step"]), d["breakdown_ms"], d["roofline"]["ms_per_launch_by_species"], d["sort_p"]["ms_per_call"])
PY
