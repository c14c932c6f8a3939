#!/bin/bash
# closing pass after the walk-code rewrite: whole GPU suite, smoke, headline bench, the other bench lines
set -u
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r02d_gpu_suite.log 2>&1; echo "suite rc=$?" >> $O/r02d_gpu_suite.log; tail -n 3 $O/r02d_gpu_suite.log
python __graft_entry__.py smoke > $O/r02d_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 $O/r02d_smoke.log
python bench.py --steps 10 --warmup 3 > $O/r02d_bench_random1m.json 2> $O/r02d_bench.err; python -c "import json;d=json.load(open('$O/r02d_bench_random1m.json'));print('random1m',d['value'],d['e2e']['value'],d['config'].get('cut_hash'),d['config']['other_inputs']['torus100_lp_replay'])"
for w in torus100 torus3d20 random20k; do
  python bench.py --workload $w --steps 20 --no-extra --no-cpu-baseline > $O/r02d_bench_$w.json 2>> $O/r02d_bench.err; python -c "import json;d=json.load(open('$O/r02d_bench_$w.json'));print('$w',d['value'],d['e2e']['value'])"
done
for w in torus100 torus3d20; do
  python bench.py --workload $w --xstar replay --steps 30 --no-extra --no-cpu-baseline > $O/r02d_bench_${w}_lp_replay.json 2>> $O/r02d_bench.err; python -c "import json;d=json.load(open('$O/r02d_bench_${w}_lp_replay.json'));print('$w replay',d['value'],d['e2e']['value'])"
done
for w in torus100 random20k torus3d20; do timeout 300 python scripts/profile_round.py --workload $w --xstar lp_like --rounds 3 2>&1 | tail -1 | python -c "import sys,ast;d=ast.literal_eval(sys.stdin.read());print('$w lp_like ms_search',d['ms_search'],d['tier_sources'])"; done
