#!/bin/bash
# Round-2 measurement pass (one B200, under gpurun): every bench line that DESIGN.md / profiles/README.md quote.
set -u
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02_gputest_final.log 2>&1; echo "pytest rc $?" >> $O/r02_gputest_final.log
python __graft_entry__.py smoke > $O/r02_smoke_final.log 2>&1
python bench.py --steps 10 --warmup 3 > $O/r02_bench_random1m.json 2> $O/r02_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_bench_reference_arm_random1m.json 2>> $O/r02_bench.err
python bench.py --mode 1 --steps 1 --no-extra --no-cpu-baseline > $O/r02_bench_random1m_modeV.json 2>> $O/r02_bench.err
for w in torus100 torus3d20 random20k; do
  python bench.py --workload $w --steps 20 --no-extra > $O/r02_bench_$w.json 2>> $O/r02_bench.err
  python bench.py --workload $w --mode 1 --steps 5 --no-extra --no-cpu-baseline > $O/r02_bench_${w}_modeV.json 2>> $O/r02_bench.err
  python bench.py --impl reference --workload $w --steps 3 --warmup 1 > $O/r02_bench_reference_arm_$w.json 2>> $O/r02_bench.err
  timeout 300 python scripts/profile_round.py --workload $w --xstar lp_like --rounds 3 > $O/r02_prof_${w}_lplike.log 2>&1
done
for w in torus100 torus3d20; do
  if [ -f tests/golden/lpseq_$( [ $w = torus100 ] && echo torus2d_100 || echo torus3d_20 ).npz ]; then
    python bench.py --workload $w --xstar replay --steps 30 --no-extra > $O/r02_bench_${w}_lp_replay.json 2>> $O/r02_bench.err
    timeout 600 python bench.py --workload $w --xstar replay --mode 1 --steps 10 --no-extra --no-cpu-baseline > $O/r02_bench_${w}_lp_replay_modeV.json 2>> $O/r02_bench.err
    python bench.py --impl reference --workload $w --xstar replay --steps 30 --warmup 1 > $O/r02_bench_reference_arm_${w}_lp_replay.json 2>> $O/r02_bench.err
  fi
done
python bench.py --workload dense --steps 10 > $O/r02_bench_dense.json 2>> $O/r02_bench.err
python bench.py --impl reference --workload dense --steps 2 --warmup 1 > $O/r02_bench_reference_arm_dense.json 2>> $O/r02_bench.err
tail -3 $O/r02_gputest_final.log
