#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none -k regex:k_wf_trace -s 6 -c 2 -o gpurun_out/s3b_wf_trace_grid4m -f python scripts/profile_trace.py wavefront 64 1920 1080 grid4m > gpurun_out/s3b_ncu3.log 2>&1; echo "ncu grid4m rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:k_wf_shade -s 8 -c 2 -o gpurun_out/s3b_wf_shade_grid4m -f python scripts/profile_trace.py wavefront 64 1920 1080 grid4m > gpurun_out/s3b_ncu3b.log 2>&1; echo "ncu grid4m shade rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 6 -c 2 -o gpurun_out/s3b_wf_trace_many303 -f python scripts/profile_trace.py wavefront 64 1920 1080 many303 > gpurun_out/s3b_ncu4.log 2>&1; echo "ncu many303 rc=$?"
