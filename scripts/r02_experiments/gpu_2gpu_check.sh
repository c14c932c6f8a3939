#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/r02b_multi_gpu_tests.log 2>&1; echo "rc=$?"; tail -n 2 gpurun_out/r02b_multi_gpu_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/r02b_bench_random1m_2gpu.json 2> gpurun_out/r02b_bench_2gpu.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/r02b_bench_random1m_2gpu.json; grep -o '"cut_hash": "[0-9a-f]*"' gpurun_out/r02b_bench_random1m_2gpu.json
