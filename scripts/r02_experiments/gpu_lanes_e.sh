#!/bin/bash
# fifth GPU pass: random-sector microbenchmark; pairs engine with the predecessor in the queue entry (parity + timing against
# the predecessor-array build); ncu of the lanes search kernel v3
mkdir -p gpurun_out
nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/sector_bw scripts/sector_bw.cu && timeout 300 /tmp/sector_bw > gpurun_out/r02_sector_bw.log 2>&1; cat gpurun_out/r02_sector_bw.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "raw_walks or tier1 or tiers_agree or history or near_cut" > gpurun_out/pairs_e_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/pairs_e_tests.log
timeout 600 python scripts/exp_lanes.py --rounds 4 --settings "0:0:0" > gpurun_out/pairs_e_new.log 2>&1; grep "round" gpurun_out/pairs_e_new.log
SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_predarr.so timeout 600 python scripts/exp_lanes.py --rounds 4 --settings "0:0:0" > gpurun_out/pairs_e_old.log 2>&1; grep "round" gpurun_out/pairs_e_old.log
P="python scripts/exp_lanes.py --rounds 2 --settings 1:98304:2:2048:14"
export SMSGPU_LIB=$PWD/sms_b200/csrc/libsmsgpu_g4.so
timeout 600 $P > gpurun_out/lanes_e_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:ocwLanesSearchKernel -s 12 -c 1 -o gpurun_out/r02_lanes_search_v3 $P > gpurun_out/lanes_e_ncu.log 2>&1
tail -n 2 gpurun_out/lanes_e_plain.log; tail -n 2 gpurun_out/lanes_e_ncu.log
