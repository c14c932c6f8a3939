#!/bin/bash
# Round-2 closing pass on one B200 (tree with the lanes engine opt-in, predecessor in the queue entry, K6 scan rewrite):
# whole GPU suite, smoke, headline bench + launch list, dense bench, reference arm, A/B of the plain-load variant.
set -u
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r02b_gpu_suite.log 2>&1; echo "suite rc=$?" >> $O/r02b_gpu_suite.log; tail -n 3 $O/r02b_gpu_suite.log
python __graft_entry__.py smoke > $O/r02b_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $O/r02b_smoke.log
python bench.py --steps 10 --warmup 3 > $O/r02b_bench_random1m.json 2> $O/r02b_bench.err; cat $O/r02b_bench_random1m.json | cut -c1-600
python bench.py --workload dense --steps 10 > $O/r02b_bench_dense.json 2>> $O/r02b_bench.err
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra"
$B > $O/r02b_ncu_bench_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02b_launches_bench_random1m.csv $B > $O/r02b_ncu_bench.log 2>&1
for i in 1 2; do
python scripts/exp_lanes.py --rounds 3 --settings "0:0:0" 2>&1 | grep "round 2" | cut -c1-110
SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_plainld.so python scripts/exp_lanes.py --rounds 3 --settings "0:0:0" 2>&1 | grep "round 2" | cut -c1-110 | sed 's/^/PLAINLD /'
done
python bench.py --impl reference --steps 2 --warmup 1 > $O/r02b_bench_reference_arm_random1m.json 2>> $O/r02b_bench.err; cut -c1-300 $O/r02b_bench_reference_arm_random1m.json
