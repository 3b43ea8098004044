#!/bin/bash
# ncu captures after the persistent shade stage / fused small build (one gpurun call). Every command runs plain first.
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
python tools/exp_wf_profile.py 2237 4 5 > gpurun_out/ncu_plain_wf.log 2>&1 &&
$NCU -k regex:rd_wf_shade -s 0 -c 1 -f -o gpurun_out/ncu_r2h_wf_shade python tools/exp_wf_profile.py 2237 4 5 > gpurun_out/ncu_run_wf2.log 2>&1
$NCU -k regex:rd_wf_shade -s 3 -c 1 -f -o gpurun_out/ncu_r2h_wf_shade_bounce3 python tools/exp_wf_profile.py 2237 4 5 > gpurun_out/ncu_run_wf3.log 2>&1
python tools/exp_smallbuilds.py 16 once > gpurun_out/ncu_plain_sb.log 2>&1 &&
$NCU -k regex:rd_k_small_build_fused -s 40 -c 1 -f -o gpurun_out/ncu_r2h_small_build_fused python tools/exp_smallbuilds.py 16 once > gpurun_out/ncu_run_sb.log 2>&1
python bench.py --steps 2 --warmup 3 --extra C3 --no-cpu-baseline > gpurun_out/launches_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r2h.csv python bench.py --steps 2 --warmup 3 --extra C3 --no-cpu-baseline > gpurun_out/launches_run.log 2>&1
ls -la gpurun_out/*.ncu-rep
