#!/bin/bash
# K6 scan rewrite: parity, bench, ncu; DRAM bytes of the random-sector microbenchmark under ncu
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_small_cycles.py tests/test_zz_no_alloc.py tests/test_host_cpp.py -x -q -m gpu > gpurun_out/k6_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/k6_tests.log
D="python bench.py --workload dense --steps 10 --no-cpu-baseline"
$D > gpurun_out/r02_bench_dense_v2.json 2> gpurun_out/k6_bench.err; cat gpurun_out/r02_bench_dense_v2.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:smallCycleScan -s 6 -c 1 -o gpurun_out/r02_smallcycle_scan_v2 python bench.py --workload dense --steps 2 --no-cpu-baseline > gpurun_out/k6_ncu.log 2>&1; tail -n 2 gpurun_out/k6_ncu.log
nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/sector_bw scripts/sector_bw.cu && timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none --csv --log-file gpurun_out/r02_sector_bw_ncu.csv -c 12 /tmp/sector_bw > gpurun_out/sector_bw_ncu.log 2>&1; tail -n 3 gpurun_out/sector_bw_ncu.log
