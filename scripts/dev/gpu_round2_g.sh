#!/bin/bash
mkdir -p gpurun_out/r2g
O=gpurun_out/r2g
for cfg in lm_ddp cifar_ddp; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_$cfg.csv python bench.py --config $cfg --steps 2 --warmup 3 --no-graph --no-e2e > $O/ncu_$cfg.log 2>&1
python scripts/launch_summary.py $O/launches_$cfg.csv 45 adamw > $O/launches_${cfg}_summary.txt 2>&1
done
cat $O/launches_lm_ddp_summary.txt $O/launches_cifar_ddp_summary.txt
