#!/bin/bash
mkdir -p gpurun_out
SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_wide.so timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "raw_walks or tier1 or history" > gpurun_out/wide_tests.log 2>&1; echo "tests rc=$?"; tail -n 2 gpurun_out/wide_tests.log
for i in 1 2; do
timeout 600 python scripts/exp_lanes.py --rounds 3 --settings "0:0:0" 2>&1 | grep "round [12]"
SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_wide.so timeout 600 python scripts/exp_lanes.py --rounds 3 --settings "0:0:0" 2>&1 | grep "round [12]" | sed 's/^/WIDE /'
done
