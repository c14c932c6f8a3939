#!/bin/bash
# pairs engine: 7 / 8 CTAs per SM (72 / 64 registers, 1536 / 1024 buckets so that the shared memory fits)
for v in "" b1536 c7 c8; do
  if [ -n "$v" ]; then export SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_$v.so; fi
  SMSGPU_DEBUG_OVF=1 python scripts/exp_lanes.py --rounds 3 --settings "0:0:0" 2>&1 | grep "round 2\|tier 1: " | tail -2 | cut -c1-230 | sed "s/^/[$v] /"
done
