#!/bin/bash
# DRAM cost of one random access: plain timing, then ncu DRAM/L2 byte counters per launch
set -x
mkdir -p gpurun_out
tools/randbench2.bin 32 > gpurun_out/randbench2_plain32.txt 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_write.sum,lts__t_sectors_srcunit_tex_lookup_miss.sum \
  --clock-control none --csv --log-file gpurun_out/randbench2_ncu32.csv tools/randbench2.bin 32 > gpurun_out/randbench2_ncu32.log 2>&1
tools/randbench2.bin 128 > gpurun_out/randbench2_plain128.txt 2>&1
cat gpurun_out/randbench2_plain32.txt
