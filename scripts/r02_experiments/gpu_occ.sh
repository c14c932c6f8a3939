#!/bin/bash
# pairs engine: fewer CTAs per SM with more registers, with and without the neighbour's row carried by the arc record
for v in "" c5 c5row c4row; do
  if [ -n "$v" ]; then export SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_$v.so; fi
  python scripts/exp_lanes.py --rounds 3 --settings "0:0:0" 2>&1 | grep "round 2" | cut -c1-120 | sed "s/^/[$v] /"
done
