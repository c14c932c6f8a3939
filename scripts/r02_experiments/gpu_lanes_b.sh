#!/bin/bash
# second GPU pass of the lanes engine: what breaks when tiles are reused (tile sets = 1 / 2, fences), ncu of the search kernel
mkdir -p gpurun_out
T="python -m pytest tests/test_lanes_gpu.py -x -q -m gpu -k batches_and_sets"
timeout 300 $T > gpurun_out/lanes_b_default.log 2>&1; echo "default rc=$?"; grep -c "lanes finish: tile" gpurun_out/lanes_b_default.log; grep "lanes finish: tile" gpurun_out/lanes_b_default.log | head -5
SMSGPU_LANES_SETS=1 timeout 300 $T > gpurun_out/lanes_b_sets1.log 2>&1; echo "sets=1 rc=$?"; grep -c "lanes finish: tile" gpurun_out/lanes_b_sets1.log
SMSGPU_LANES_FENCE=1 timeout 300 $T > gpurun_out/lanes_b_fence.log 2>&1; echo "fence rc=$?"; grep -c "lanes finish: tile" gpurun_out/lanes_b_fence.log
tail -n 5 gpurun_out/lanes_b_sets1.log gpurun_out/lanes_b_fence.log
P="python scripts/exp_lanes.py --rounds 2 --sources 81920 --settings 1:40960:2"
timeout 600 $P > gpurun_out/lanes_b_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:ocwLanesSearchKernel -s 2 -c 1 -o gpurun_out/r02_lanes_search_v1 $P > gpurun_out/lanes_b_ncu.log 2>&1
tail -n 4 gpurun_out/lanes_b_plain.log; tail -n 3 gpurun_out/lanes_b_ncu.log; ls -la gpurun_out/*.ncu-rep
