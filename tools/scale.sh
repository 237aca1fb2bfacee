# usage: tools/scale.sh N  -> gpurun_out/scale_N.json (full-size weak-scaling bench on N GPUs, no e2e / cpu / yee legs)
N=$1; O=gpurun_out; mkdir -p $O
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu --no-e2e --no-yee --breakdown 4 2>$O/scale_$N.err | tail -n 1 > $O/scale_$N.json
python - <<PY
import json
d=json.loads(open("gpurun_out/scale_$N.json").read().strip().splitlines()[-1])
print("N=$N Gpush/s %.2f ms/step %.3f"%(d["value"]/1e9,d["ms_per_step"]), d["breakdown_ms"], d["config"]["parallelism"][:120], d["clocks"])
PY
