mkdir -p gpurun_out
source <(sed -n '/^run() {/,/^}/p' tools/run_l.sh | sed 's/r02_m_/r02_n_/g')
run c2k33 --workload C2 --k 33 --no-parity
run c2k55 --workload C2 --k 55 --no-parity
run k77m15 --workload C4 --k 77 --m 15
run k77m27 --workload C4 --k 77 --m 27 --no-parity
run k99m15 --workload C4 --k 99 --m 15 --no-parity
run k99m27 --workload C4 --k 99 --m 27
