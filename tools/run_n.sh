# usage: bash tools/run_n.sh N  -- multi-GPU experiments on one box (gpurun --gpus N)
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for X in ${XMODES:-pull copy}; do
  echo "== exchange $X"
  timeout 600 $TR bench.py --gpus $N --exchange $X --configs none --no-e2e --steps 5 --warmup 3 > gpurun_out/r02_n${N}_$X.json 2> gpurun_out/r02_n${N}_$X.err
  echo "rc=$?"; tail -2 gpurun_out/r02_n${N}_$X.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r02_n${N}_$X.json').read().strip().splitlines()[-1])
    print(round(d['value']/1e9,2), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['phases_ms_per_step'].items() if v}, d['step_wall_ms'])
    print('parity', d['parity_check'], 'exchange', d.get('exchange'), 'nccl', d.get('exchange_nccl'))
except Exception as e: print('no line', e)
PY
done
echo "== pcie probe"
timeout 300 $TR tools/ubench/pcie_probe.py > gpurun_out/r02_pcie_n${N}.txt 2>&1; head -c 3000 gpurun_out/r02_pcie_n${N}.txt
echo "== full default (pull) with e2e and configs"
timeout 900 $TR bench.py --gpus $N > gpurun_out/r02_n${N}_full.json 2> gpurun_out/r02_n${N}_full.err
echo "rc=$?"; tail -2 gpurun_out/r02_n${N}_full.err
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r02_n${N}_full.json').read().strip().splitlines()[-1])
    print(round(d['value']/1e9,2), round(d['ms_per_step'],2))
    print('e2e', {k:v for k,v in d['e2e'].items() if k not in ('api',)})
    for k,v in d.get('configs',{}).items(): print(k, v if 'error' in v else (round(v['value']/1e9,2), round(v['ms_per_step'],1), {a:round(b,1) for a,b in v['phases_ms_per_step'].items() if b}, v['split_passes']))
except Exception as e: print('no line', e)
PY
