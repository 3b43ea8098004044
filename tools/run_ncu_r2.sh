#!/bin/bash
# ncu captures of round 2 (one gpurun call): the inline raygen on C2 and on the 10 M-triangle field, the first-bounce trace and
# shade stages of C3, and the launch list of a short bench run. Every command runs plain first.
mkdir -p gpurun_out
export RD_EXP_VARIANTS='[{"flags":""}]'
NCU="ncu --set full --clock-control none --import-source on"
for n in 708 2237; do
  python tools/exp_variants.py $n > gpurun_out/ncu_plain_$n.log 2>&1 &&
  $NCU -k regex:^raygen -s 10 -c 1 -f -o gpurun_out/ncu_r2_primary_$n python tools/exp_variants.py $n > gpurun_out/ncu_run_$n.log 2>&1
done
python tools/exp_wf_profile.py 2237 4 5 > gpurun_out/ncu_plain_wf.log 2>&1 &&
$NCU -k regex:rd_k_wf_trace_dyn -s 0 -c 1 -f -o gpurun_out/ncu_r2_wf_trace_dyn python tools/exp_wf_profile.py 2237 4 5 > gpurun_out/ncu_run_wf1.log 2>&1
$NCU -k regex:rd_wf_shade -s 0 -c 1 -f -o gpurun_out/ncu_r2_wf_shade python tools/exp_wf_profile.py 2237 4 5 > gpurun_out/ncu_run_wf2.log 2>&1
$NCU -k regex:rd_k_wf_trace$ -s 0 -c 1 -f -o gpurun_out/ncu_r2_wf_trace_primary python tools/exp_wf_profile.py 2237 4 5 > gpurun_out/ncu_run_wf3.log 2>&1
python bench.py --steps 2 --warmup 3 --extra C3 --no-cpu-baseline > gpurun_out/launches_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r2.csv python bench.py --steps 2 --warmup 3 --extra C3 --no-cpu-baseline > gpurun_out/launches_run.log 2>&1
ls -la gpurun_out/*.ncu-rep
