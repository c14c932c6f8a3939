#!/bin/bash
# Round-2 profiling pass (run under gpurun on one B200): every ncu command right behind the same command run plain.
set -u
O=gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra"
$B > $O/r02_ncu_bench_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches_bench_random1m.csv $B > $O/r02_ncu_bench.log 2>&1
P="python scripts/profile_round.py --sources 262144 --rounds 2 --start-tier 1"
$P > $O/r02_ncu_pairs_plain.log 2>&1 && timeout 1500 ncu --set full --clock-control none --import-source on -k regex:ocwSearchPairsKernel -s 1 -c 1 -o $O/r02_pairs_tier1_random1m_256k $P > $O/r02_ncu_pairs.log 2>&1
R="python scripts/profile_round.py --workload random20k --xstar lp_like --sources 4000 --rounds 2 --start-tier 3"
$R > $O/r02_ncu_runs_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:ocwSearchRunsKernel -s 1 -c 1 -o $O/r02_runs_tier3_random20k_lplike_4k $R > $O/r02_ncu_runs.log 2>&1
D="python bench.py --workload dense --steps 2 --no-cpu-baseline"
$D > $O/r02_ncu_dense_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:smallCycle -s 6 -c 2 -o $O/r02_smallcycle_dense_k500 $D > $O/r02_ncu_dense.log 2>&1
ls -la $O/*.ncu-rep
