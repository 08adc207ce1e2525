#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/probe_reuse.py > gpurun_out/reuse_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_requests_srcunit_tex_op_read.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:k_line_reuse --csv --log-file gpurun_out/reuse_ncu.csv python tools/probe_reuse.py > gpurun_out/reuse_ncu.log 2>&1
echo rc=$?; cat gpurun_out/reuse_plain.log
