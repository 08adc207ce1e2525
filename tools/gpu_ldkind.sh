#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/probe_ldkind.py > gpurun_out/ldkind_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,lts__t_requests_srcunit_tex_op_read.sum --clock-control none -k regex:k_gather --csv --log-file gpurun_out/ldkind_ncu.csv python tools/probe_ldkind.py > gpurun_out/ldkind_ncu.log 2>&1
echo rc=$?; cat gpurun_out/ldkind_plain.log
