mkdir -p gpurun_out
export PYTHONPATH=$PWD
( timeout 120 python -m pytest tests/test_merge.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02_merge_gpu3.log 2>&1; echo "merge+noalloc rc=$?"; tail -3 gpurun_out/r02_merge_gpu3.log )
timeout 60 python tools/merge_probe.py 1000 3
