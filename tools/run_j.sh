mkdir -p gpurun_out
( timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/r02_gputest_j.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_gputest_j.log )
( timeout 420 python bench.py --no-e2e --no-cpu-baseline --no-lookup > gpurun_out/r02_bench_j.json 2> gpurun_out/r02_bench_j.err; echo "bench rc=$?" ); tail -2 gpurun_out/r02_bench_j.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_j.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], {k:round(v,2) for k,v in d['phases_ms_per_step'].items() if v}, d['reference_pin'])
for k,v in d.get('configs',{}).items(): print(k, v if 'error' in v else (round(v['value']/1e9,2), round(v['ms_per_step'],1), round(v['roofline_frac'],3), {a:round(b,1) for a,b in v['phases_ms_per_step'].items() if b}, v['split_passes'], v['reference_pin_ok']))
PY
