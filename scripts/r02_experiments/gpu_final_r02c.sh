#!/bin/bash
# last pass: smoke (tier 0, pairs, lanes), bench lines of configs 2-4 and the LP replays on the final tree, ncu of the pairs kernel
set -u
O=gpurun_out
mkdir -p $O
python __graft_entry__.py smoke > $O/r02c_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 $O/r02c_smoke.log
for w in torus100 torus3d20 random20k; do
  python bench.py --workload $w --steps 20 --no-extra --no-cpu-baseline > $O/r02c_bench_$w.json 2>> $O/r02c_bench.err; python -c "import json;d=json.load(open('$O/r02c_bench_$w.json'));print('$w',d['value'],d['e2e']['value'])"
done
for w in torus100 torus3d20; do
  python bench.py --workload $w --xstar replay --steps 30 --no-extra --no-cpu-baseline > $O/r02c_bench_${w}_lp_replay.json 2>> $O/r02c_bench.err; python -c "import json;d=json.load(open('$O/r02c_bench_${w}_lp_replay.json'));print('$w replay',d['value'],d['e2e']['value'])"
done
P="python scripts/profile_round.py --sources 262144 --rounds 2 --start-tier 1"
$P > $O/r02c_ncu_pairs_plain.log 2>&1 && timeout 1500 ncu --set full --clock-control none --import-source on -k regex:ocwSearchPairsKernel -s 1 -c 1 -o $O/r02c_pairs_tier1_random1m_256k $P > $O/r02c_ncu_pairs.log 2>&1; tail -n 2 $O/r02c_ncu_pairs.log
