# ncu captures of this round (one GPU): launch list of the default bench command, then --set full of the dominant kernels
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-lookup --no-parity --configs none"
P="--steps 1 --warmup 0 --no-e2e --no-cpu-baseline --no-lookup --no-parity --configs none --reads-per-gpu 4000000 --n-genomes 20"
for K in 33 21 77; do
  python bench.py $P --k $K > gpurun_out/r02_prof_k$K.json 2>/dev/null && \
  ncu --set full --clock-control none --import-source on -k regex:"count_kernel|refine_bulk_kernel" -c 2 -o gpurun_out/r02_full_k$K -f python bench.py $P --k $K > gpurun_out/r02_ncu_full_k$K.log 2>&1
  echo "k=$K rc=$?"
done
ncu --set full --clock-control none --import-source on -k regex:"parse_kernel" -s 1 -c 1 -o gpurun_out/r02_full_parse_k33 -f python bench.py $P --k 33 > gpurun_out/r02_ncu_full_parse.log 2>&1
echo "parse rc=$?"
ls -la gpurun_out/*.ncu-rep
