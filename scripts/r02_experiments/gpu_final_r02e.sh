#!/bin/bash
# pool scans of the runs kernel with four keys per thread in flight: whole GPU suite, smoke, LP replays
set -u
O=gpurun_out
mkdir -p $O
python bench.py --workload torus100 --xstar replay --steps 30 --no-extra --no-cpu-baseline > $O/r02e_bench_torus100_lp_replay.json 2>> $O/r02e_bench.err; python -c "import json;d=json.load(open('$O/r02e_bench_torus100_lp_replay.json'));print('torus100 replay',d['value'],d['e2e']['value'])"
python bench.py --workload torus3d20 --xstar replay --steps 30 --no-extra --no-cpu-baseline > $O/r02e_bench_torus3d20_lp_replay.json 2>> $O/r02e_bench.err; python -c "import json;d=json.load(open('$O/r02e_bench_torus3d20_lp_replay.json'));print('torus3d20 replay',d['value'],d['e2e']['value'])"
timeout 300 python scripts/profile_round.py --workload random20k --xstar lp_like --rounds 3 2>&1 | tail -1 | python -c "import sys,ast;d=ast.literal_eval(sys.stdin.read());print('random20k lp_like ms_search',d['ms_search'],d['tier_sources'])"
timeout 1200 python -m pytest tests -x -q -m gpu > $O/r02e_gpu_suite.log 2>&1; echo "suite rc=$?" >> $O/r02e_gpu_suite.log; tail -n 3 $O/r02e_gpu_suite.log
python __graft_entry__.py smoke > $O/r02e_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 $O/r02e_smoke.log
