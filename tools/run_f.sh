mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_pull_exchange.py tests/test_gpu_configs.py -x -q -m gpu > gpurun_out/r02_gputest_f.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_gputest_f.log )
Q="--no-e2e --no-cpu-baseline --no-lookup --no-parity --configs none --steps 4 --warmup 3"
for V in "KC_REFINE_BULK=0" "KC_REFINE_BULK=1" "KC_REFINE_BULK=1 KC_REFINE_BULK_SHAPE=1"; do
  echo "$V"
  env $V timeout 300 python bench.py $Q 2>>gpurun_out/r02_f.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['checksum'], d['ms_per_step'], d['phases_ms_per_step']['refine_kernel'], d['phases_ms_per_step']['insert_kernel'])"
done
tail -3 gpurun_out/r02_f.err
