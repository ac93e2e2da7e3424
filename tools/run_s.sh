mkdir -p gpurun_out
export PYTHONPATH=$PWD
( timeout 100 python -m pytest tests/test_merge.py -x -q -m gpu -k "feed_the_stage" > gpurun_out/r02_merge_gpu5.log 2>&1; echo "feed rc=$?"; tail -12 gpurun_out/r02_merge_gpu5.log )
