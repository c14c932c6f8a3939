#!/bin/bash
mkdir -p gpurun_out
nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/sector_flavors scripts/sector_flavors.cu || exit 1
/tmp/sector_flavors > gpurun_out/r02_sector_flavors.log 2>&1; cat gpurun_out/r02_sector_flavors.log
M=dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,gpu__time_duration.sum
timeout 600 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/r02_sector_flavors_ncu.csv /tmp/sector_flavors > /dev/null 2>&1
timeout 600 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/r02_sector_flavors_ncu_g32.csv /tmp/sector_flavors 32 > gpurun_out/r02_sector_flavors_g32.log 2>&1; head -2 gpurun_out/r02_sector_flavors_g32.log
