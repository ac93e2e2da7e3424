# experiments: pre-split factor (KC_PRESPLIT) and sample cap (KC_SAMPLE_MAX_FRAC) on the long-k configs; ncu of the k = 77 count kernel
mkdir -p gpurun_out
run() {  # label presplit frac extra-args...
  local label=$1 f=$2 fr=$3; shift 3
  KC_PRESPLIT=$f KC_SAMPLE_MAX_FRAC=$fr timeout 200 python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu-baseline --no-lookup --configs none "$@" \
    > gpurun_out/r02_k_$label.json 2> gpurun_out/r02_k_$label.err || { echo "$label FAILED rc=$?"; tail -3 gpurun_out/r02_k_$label.err; }
  python - "$label" "$f" "$fr" <<'PY'
import json, sys
label = sys.argv[1]
try:
    d = json.loads(open(f'gpurun_out/r02_k_{label}.json').read().strip().splitlines()[-1])
    ph = {k: round(v, 1) for k, v in d['phases_ms_per_step'].items() if v}
    pc = d.get('parity_check')
    print(label, 'presplit', sys.argv[2], 'frac', sys.argv[3], '| step', round(d['ms_per_step'], 1), ph, 'split', d['table']['split_passes'],
          'hbm GB', round(d['hbm_bytes_in_use'] / 1e9, 1), 'parity', None if pc is None else pc['ok'], 'pin', (d.get('reference_pin') or {}).get('ok'),
          'stable', d['result_stability']['identical'], 'chk', d['checksum'])
except Exception as e:
    print(label, 'no result', e)
PY
}
run k77m15_base 0 1.0 --workload C4 --k 77 --m 15 --no-parity
run k77m15_f10 1.0 0.2 --workload C4 --k 77 --m 15
run k77m15_f20 2.0 0.2 --workload C4 --k 77 --m 15 --no-parity
run k77m15_f05 0.5 0.2 --workload C4 --k 77 --m 15 --no-parity
run k99m27_f10 1.0 0.2 --workload C4 --k 99 --m 27
run k99m27_f20 2.0 0.1 --workload C4 --k 99 --m 27 --no-parity
run c2k55_f10 1.0 1.0 --workload C2 --k 55
run c2k33_f10 1.0 1.0 --workload C2 --k 33
# ncu: count kernel of k = 77 with the two-stage split passes (4 M reads over 20 genomes keep the coverage of the full workload)
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"count_kernel" -c 1 -o gpurun_out/r02_full_k77_split2 -f \
  python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --no-lookup --no-parity --configs none --workload C4 --reads-per-gpu 4000000 --n-genomes 20 --k 77 --m 15 \
  > gpurun_out/r02_ncu_k77.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep 2>/dev/null
