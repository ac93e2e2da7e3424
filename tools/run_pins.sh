mkdir -p gpurun_out
cp tests/golden/full_config_pins.json gpurun_out/full_config_pins.json
for A in "C2 21 20000000" "C2 55 20000000" "C5 21 20000000" "C1 21 1000000"; do set -- $A; timeout 600 python tools/pin_full_config.py --workload $1 --k $2 --reads $3 | cut -c1-400; done
( timeout 900 python -m pytest tests/test_io.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r02_gputest_h.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_gputest_h.log )
