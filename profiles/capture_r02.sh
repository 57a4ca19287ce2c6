#!/bin/bash
# ncu evidence of round 2 (run on the GPU box from the repo root; outputs under gpurun_out/).
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-long-legs --no-cpu-baseline"
$CMD > gpurun_out/r02_plain.json 2> gpurun_out/r02_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_cfg3.csv $CMD > gpurun_out/r02_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fused_tmem_pipe -s 2 -c 1 -o gpurun_out/r02_fused_cfg3_full -f $CMD --no-e2e > gpurun_out/r02_ncu_f.log 2>&1
python profiles/microbench/cfg2_run.py > gpurun_out/r02_cfg2_run.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:fused_cta -s 1 -c 1 -o gpurun_out/r02_fused_cta_cfg2 -f python profiles/microbench/cfg2_run.py > gpurun_out/r02_ncu_c.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02_launches_cfg2.csv python profiles/microbench/cfg2_run.py > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:thomas_steps -s 3 -c 1 -o gpurun_out/r02_thomas_v1 -f $CMD --no-e2e > gpurun_out/r02_ncu_t.log 2>&1
ls -la gpurun_out/*.ncu-rep
