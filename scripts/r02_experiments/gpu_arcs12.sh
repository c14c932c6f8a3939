#!/bin/bash
SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_arcs12.so timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tier1_engines or history" 2>&1 | tail -2
for v in "" arcs12 "" arcs12; do
  if [ -n "$v" ]; then export SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_$v.so; else unset SMSGPU_LIB; fi
  python scripts/exp_lanes.py --rounds 3 --settings "0:0:0" 2>&1 | grep "round 2" | cut -c1-120 | sed "s/^/[$v] /"
done
