mkdir -p gpurun_out
export PYTHONPATH=$PWD
( timeout 150 python -m pytest tests/test_merge.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02_merge_gpu4.log 2>&1; echo "merge+noalloc rc=$?"; tail -3 gpurun_out/r02_merge_gpu4.log )
