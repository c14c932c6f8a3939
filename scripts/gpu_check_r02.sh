#!/bin/bash
# Round-2 confirmation pass on one B200: the no-allocation entry point's test alone (fresh process), the whole GPU suite
# (which ends with that test again), smoke, the default bench line.  Outputs under gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/na_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/na_alone.log
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/gpu_suite.log 2>&1; echo "suite rc=$?" >> gpurun_out/gpu_suite.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err
tail -n 4 gpurun_out/na_alone.log gpurun_out/gpu_suite.log gpurun_out/smoke.log; cat gpurun_out/bench_final.json
