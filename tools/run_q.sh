mkdir -p gpurun_out
( timeout 330 python bench.py > gpurun_out/r02_bench_final2.json 2> gpurun_out/r02_bench_final2.err; echo "bench rc=$?" ); tail -2 gpurun_out/r02_bench_final2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_final2.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], {k:round(v,2) for k,v in d['phases_ms_per_step'].items() if v}, d['reference_pin'], d['roofline']['frac'])
print('e2e', {k:v for k,v in d['e2e'].items() if k in ('value','ms_per_step','h2d_bytes_per_step','d2h_bytes_per_step')})
print('cpu', d.get('cpu_baseline',{}).get('value'), 'merge', d.get('merge_pairs'))
for k,v in d.get('configs',{}).items(): print(k, v if 'error' in v else (round(v['value']/1e9,2), round(v['ms_per_step'],1), round(v['roofline_frac'],3), v['split_passes'], v['reference_pin_ok']))
PY
