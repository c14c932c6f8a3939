#!/bin/bash
# third GPU pass: lanes engine v2 (phased groups of arcs, atomics) -- parity tests, then timings incl. the L2 fetch granularity
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_lanes_gpu.py -x -q -m gpu > gpurun_out/lanes_c_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/lanes_c_tests.log
SMSGPU_DEBUG_OVF=1 timeout 900 python scripts/exp_lanes.py --rounds 2 --settings "0:0:0:0:0:0:32,1:40960:2,1:40960:2:0:0:0:32,1:81920:2:1024:13,1:81920:2:1024:13:0:32" > gpurun_out/lanes_exp_c.log 2>&1; echo "rc=$?"
grep -v "^\[smsgpu\] tier" gpurun_out/lanes_exp_c.log | tail -n 30
SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_g4.so timeout 600 python scripts/exp_lanes.py --rounds 2 --settings "1:40960:2,1:81920:2:1024:13" > gpurun_out/lanes_exp_c_g4.log 2>&1; echo "g4 rc=$?"
grep "round 1" gpurun_out/lanes_exp_c_g4.log
