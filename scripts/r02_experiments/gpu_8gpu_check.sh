#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/r02b_bench_random1m_8gpu.json 2> gpurun_out/r02b_bench_8gpu.err; echo "bench rc=$?"; grep '^{' gpurun_out/r02b_bench_random1m_8gpu.json | cut -c1-200; grep -o '"cut_hash": "[0-9a-f]*"\|"e2e": {[^}]*}' gpurun_out/r02b_bench_random1m_8gpu.json
