mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_configs.py -x -q -m gpu > gpurun_out/r02_gputest_e.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_gputest_e.log )
for i in 1 2 3; do timeout 300 python bench.py --no-e2e --no-cpu-baseline --no-lookup --configs none 2>>gpurun_out/r02_e.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['checksum'], d['ms_per_step'], d['step_wall_ms'], d['parity_check']['ok'], d['phases_ms_per_step']['insert_kernel'], d['table']['split_passes'])"; done
tail -3 gpurun_out/r02_e.err
