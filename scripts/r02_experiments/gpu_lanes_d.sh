#!/bin/bash
# fourth GPU pass: lanes engine v3 (compact tile: dense queue entries, one sector per bucket) -- parity tests, then timings of
# register-capped builds at growing numbers of searches in flight
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_lanes_gpu.py -x -q -m gpu > gpurun_out/lanes_d_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/lanes_d_tests.log
for v in g4 g4b10 g2b16 g8b6; do
  SMSGPU_DEBUG_OVF=1 SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_$v.so timeout 600 python scripts/exp_lanes.py --rounds 2 --settings "1:65536:2:2048:13,1:131072:2:2048:13,1:98304:2:2048:14" > gpurun_out/lanes_exp_d_$v.log 2>&1; echo "$v rc=$?"
  grep "lanes engine: . sets\|round 1\|left for" gpurun_out/lanes_exp_d_$v.log | grep -v "^\[smsgpu\] tier"
done
