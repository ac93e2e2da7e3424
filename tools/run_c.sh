mkdir -p gpurun_out
free -g | head -2; nproc
( timeout 1200 python -m pytest tests/test_gpu_configs.py -x -q -m gpu > gpurun_out/r02_gputest_configs.log 2>&1; echo "pytest configs rc=$?"; tail -15 gpurun_out/r02_gputest_configs.log )
Q="--no-e2e --no-cpu-baseline --no-lookup --no-parity --configs none --steps 4 --warmup 3"
for L in 0.6 0.75 0.9 1.05; do
  echo "KC_BIN_LOAD=$L"
  KC_BIN_LOAD=$L timeout 300 python bench.py $Q 2>>gpurun_out/r02_knobs.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['phases_ms_per_step'], d['table']['local_bins'], d['table']['split_passes'])"
done
( timeout 1200 python bench.py > gpurun_out/r02_bench_c.json 2> gpurun_out/r02_bench_c.err; echo "full rc=$?" )
tail -3 gpurun_out/r02_bench_c.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_c.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['phases_ms_per_step'], d['step_wall_ms'])
print('parity', d['parity_check'], d.get('reference_pin'), d['result_stability'])
print('e2e', {k:v for k,v in d['e2e'].items() if k!='api'})
print('cpu', d['cpu_baseline'])
print('roofline', d['roofline']['frac'], d['roofline']['ms_per_launch'])
for k,v in d.get('configs',{}).items(): print(k, v if 'error' in v else (round(v['value']/1e9,2), round(v['ms_per_step'],1), round(v['roofline_frac'],3), {a:round(b,1) for a,b in v['phases_ms_per_step'].items()}, v['split_passes'], v['local_bins'], v['hbm_bytes_in_use']>>20, v['table_bytes']>>20, v['filter_bytes']>>20))
PY
MEM=$(free -g | awk 'NR==2{print $7}')
if [ "$MEM" -gt 150 ]; then timeout 900 python tools/pin_full_config.py; else echo "not enough RAM for the full pin: $MEM GB"; fi
