mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_gputest_d.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02_gputest_d.log )
( timeout 300 python tools/ubench/pcie_probe.py > gpurun_out/r02_pcie_n1.txt 2>&1; echo "pcie rc=$?"; head -c 1500 gpurun_out/r02_pcie_n1.txt )
( timeout 1200 python bench.py > gpurun_out/r02_bench_d.json 2> gpurun_out/r02_bench_d.err; echo "full rc=$?" )
tail -3 gpurun_out/r02_bench_d.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_d.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['phases_ms_per_step'], d['step_wall_ms'])
print('parity', d['parity_check']['ok'], d.get('reference_pin'), d['result_stability']['identical'])
print('e2e', {k:v for k,v in d['e2e'].items() if k!='api'})
for k,v in d.get('configs',{}).items(): print(k, v if 'error' in v else (round(v['value']/1e9,2), round(v['ms_per_step'],1), round(v['roofline_frac'],3), {a:round(b,1) for a,b in v['phases_ms_per_step'].items() if b}, v['split_passes'], v['local_bins']))
PY
