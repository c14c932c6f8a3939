#!/bin/bash
# first GPU pass of the lanes engine: its parity tests, then timings on config 5
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_lanes_gpu.py tests/test_gpu_parity.py -x -q -m gpu -k "lanes or tier1_engines or tiers_agree" > gpurun_out/lanes_tests.log 2>&1; echo "rc=$?" >> gpurun_out/lanes_tests.log
tail -n 15 gpurun_out/lanes_tests.log
SMSGPU_DEBUG_OVF=1 timeout 1200 python scripts/exp_lanes.py --rounds 3 --settings "0:0:0,1:40960:2,1:81920:2:1024:13,1:81920:4:1024:13,1:49152:2:2048:13" > gpurun_out/lanes_exp_a.log 2>&1; echo "rc=$?" >> gpurun_out/lanes_exp_a.log
grep -v "^\[smsgpu\] tier" gpurun_out/lanes_exp_a.log | tail -n 40
