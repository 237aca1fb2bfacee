O=gpurun_out; mkdir -p $O
( time timeout 900 python -m pytest tests/test_multi_rank.py -m gpu -x -q ) > $O/c5_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 15 $O/c5_pytest.log
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 "$@" 2>$O/c5_err.log | tail -n 1; }
for t in put nccl; do
  VPIC_B200_TRANSPORT=$t run --cells 128x128x128 --steps 20 --warmup 4 --no-cpu --no-e2e --no-yee > $O/c5_small_$t.json
  python -c "
import json; d=json.load(open('$O/c5_small_$t.json')); print('$t 128^3/GPU', 'Gpush/s %.2f ms/step %.3f launches %d' % (d['value']/1e9, d['ms_per_step'], d['gpu_launches']))"
done
for t in put nccl; do
  VPIC_B200_TRANSPORT=$t run --cells 32x32x32 --steps 50 --warmup 5 --no-cpu --no-e2e --no-yee > $O/c5_tiny_$t.json
  python -c "
import json; d=json.load(open('$O/c5_tiny_$t.json')); print('$t 32^3/GPU', 'Gpush/s %.2f ms/step %.3f launches %d' % (d['value']/1e9, d['ms_per_step'], d['gpu_launches']))"
done
