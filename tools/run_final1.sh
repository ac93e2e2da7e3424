mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_gputest_final.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02_gputest_final.log )
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/r02_smoke.log )
( timeout 900 python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err; echo "bench rc=$?" )
tail -3 gpurun_out/r02_bench_final.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_final.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['phases_ms_per_step'], d['step_wall_ms'])
print('parity', d['parity_check']['ok'], d.get('reference_pin'), d['result_stability']['identical'], 'roofline', d['roofline']['frac'], d['roofline']['traffic'])
print('e2e', {k:v for k,v in d['e2e'].items() if k not in ('api','from_packed_reads_container')})
print('packed', d['e2e'].get('from_packed_reads_container'))
for k,v in d.get('configs',{}).items(): print(k, v if 'error' in v else (round(v['value']/1e9,2), round(v['ms_per_step'],1), round(v['roofline_frac'],3)))
PY
( timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r02_bench_reference_arm.json 2>/dev/null; echo "ref arm rc=$?"; cut -c1-300 gpurun_out/r02_bench_reference_arm.json )
