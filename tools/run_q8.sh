N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for V in 5; do
  echo "== KC_FINE_SHIFT=$V"
  KC_FINE_SHIFT=$V timeout 300 $TR bench.py --gpus $N --configs none --no-e2e --no-lookup --steps 4 --warmup 3 > gpurun_out/r02_q${N}_$V.json 2> gpurun_out/r02_q${N}_$V.err
  echo "rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r02_q${N}_$V.json').read().strip().splitlines()[-1])
    print(round(d['value']/1e9,2), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['phases_ms_per_step'].items() if v}, d['step_wall_ms'], d['parity_check']['ok'], d['checksum'])
    print('nccl', d.get('exchange_nccl'))
except Exception as e: print('no line', e)
PY
done
