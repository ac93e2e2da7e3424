mkdir -p gpurun_out
Q="--no-e2e --no-cpu-baseline --no-lookup --no-parity --configs none --steps 4 --warmup 3"
( timeout 900 python bench.py > gpurun_out/r02_bench_b.json 2> gpurun_out/r02_bench_b.err; echo "full rc=$?" )
for S in 1 12; do for OV in 0 1; do
  echo "KC_SLICES=$S KC_REFINE_OVERLAP=$OV"
  KC_SLICES=$S KC_REFINE_OVERLAP=$OV timeout 300 python bench.py $Q 2>>gpurun_out/r02_knobs.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['phases_ms_per_step'], d['step_wall_ms'])"
done; done
tail -3 gpurun_out/r02_bench_b.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_b.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['phases_ms_per_step'], d['step_wall_ms'])
print('parity', d['parity_check'])
print('e2e', {k:v for k,v in d['e2e'].items() if k!='api'})
print('roofline', d['roofline']['frac'], d['roofline']['ms_per_launch'])
for k,v in d.get('configs',{}).items(): print(k, v if 'error' in v else (round(v['value']/1e9,2), round(v['ms_per_step'],1), round(v['roofline_frac'],3), v['phases_ms_per_step'], v['split_passes'], v['hbm_bytes_in_use']>>20))
PY
