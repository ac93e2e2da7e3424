mkdir -p gpurun_out
( KC_REFINE_SMS=20 KC_DEBUG=1 timeout 900 python -m pytest tests/test_gpu_configs.py -x -q -m gpu > gpurun_out/r02_gputest_g.log 2>&1; echo "pytest (partitioned) rc=$?"; tail -4 gpurun_out/r02_gputest_g.log; grep -c "timed out" gpurun_out/r02_gputest_g.log )
Q="--no-e2e --no-cpu-baseline --no-lookup --no-parity --configs none --steps 4 --warmup 3"
for V in "KC_REFINE_SMS=20" "KC_REFINE_SMS=40"; do
  echo "$V"
  env $V KC_DEBUG=1 timeout 300 python bench.py $Q 2>>gpurun_out/r02_g.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['checksum'], d['ms_per_step'], {k:round(v,2) for k,v in d['phases_ms_per_step'].items() if v})"
done
tail -3 gpurun_out/r02_g.err
