#!/bin/bash
set -x
tools/randbench3.bin > gpurun_out/randbench3_plain.txt 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_write.sum,lts__t_sectors_srcunit_tex_op_red.sum \
  --clock-control none --csv --log-file gpurun_out/randbench3_ncu.csv tools/randbench3.bin > gpurun_out/randbench3_ncu.log 2>&1
tail -3 gpurun_out/randbench3_ncu.log
