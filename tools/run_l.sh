# adaptive pre-split threshold + sample cap at their defaults: every resident config once
mkdir -p gpurun_out
run() {  # label extra-args...
  local label=$1; shift 1
  timeout 200 python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu-baseline --no-lookup --configs none "$@" \
    > gpurun_out/r02_m_$label.json 2> gpurun_out/r02_m_$label.err || { echo "$label FAILED rc=$?"; tail -3 gpurun_out/r02_m_$label.err; }
  python - "$label" <<'PY'
import json, sys
label = sys.argv[1]
try:
    d = json.loads(open(f'gpurun_out/r02_m_{label}.json').read().strip().splitlines()[-1])
    ph = {k: round(v, 1) for k, v in d['phases_ms_per_step'].items() if v}
    pc = d.get('parity_check')
    print(label, '| step', round(d['ms_per_step'], 1), ph, 'split', d['table']['split_passes'],
          'hbm GB', round(d['hbm_bytes_in_use'] / 1e9, 1), 'parity', None if pc is None else pc['ok'], 'pin', (d.get('reference_pin') or {}).get('ok'),
          'stable', d['result_stability']['identical'], 'chk', d['checksum'])
except Exception as e:
    print(label, 'no result', e)
PY
}
run c2k33 --workload C2 --k 33
KC_PRESPLIT_ADAPT=0 run k77m15_static --workload C4 --k 77 --m 15 --no-parity
KC_PRESPLIT_ADAPT=0 run k99m27_static --workload C4 --k 99 --m 27 --no-parity
KC_PRESPLIT=0 run c2k33_off --workload C2 --k 33 --no-parity
run c2k21 --workload C2 --k 21
run c2k55 --workload C2 --k 55
run c1k21 --workload C1 --k 21
run c5k21 --workload C5 --k 21
run k77m15 --workload C4 --k 77 --m 15
run k77m27 --workload C4 --k 77 --m 27 --no-parity
run k99m15 --workload C4 --k 99 --m 15 --no-parity
run k99m27 --workload C4 --k 99 --m 27
