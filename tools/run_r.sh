mkdir -p gpurun_out
export PYTHONPATH=$PWD
timeout 40 ncu --set full --clock-control none --import-source on -k regex:"merge_kernel" -c 1 -o gpurun_out/r02_merge_kernel_final -f python tools/merge_probe.py 200 1 > gpurun_out/r02_ncu_merge.log 2>&1; echo "ncu rc=$?"
