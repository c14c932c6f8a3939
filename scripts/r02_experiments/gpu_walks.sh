#!/bin/bash
# walk code (extractCycles / simpleOddCycles) with the lanes over independent items: parity subset, LP replay breakdown, config 5
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_lp_points.py -x -q -m gpu > gpurun_out/walks_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/walks_tests.log
SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_rclk.so python scripts/exp_lp_replay.py torus2d_100 > gpurun_out/rclk3.log 2>&1; tail -2 gpurun_out/rclk3.log | head -1 | tr '|' '\n' | tail -12
python scripts/exp_lp_replay.py torus2d_100 2>&1 | tail -2 | cut -c1-200
python scripts/exp_lanes.py --rounds 3 --settings "0:0:0" 2>&1 | grep "round 2" | cut -c1-120
python bench.py --workload torus100 --xstar replay --steps 30 --no-extra --no-cpu-baseline 2>/dev/null | python -c "import json,sys;d=json.loads(sys.stdin.read());print('torus100 replay',d['value'],d['e2e']['value'])"
python bench.py --workload torus3d20 --xstar replay --steps 30 --no-extra --no-cpu-baseline 2>/dev/null | python -c "import json,sys;d=json.loads(sys.stdin.read());print('torus3d20 replay',d['value'],d['e2e']['value'])"
