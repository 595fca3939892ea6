#!/bin/bash
mkdir -p gpurun_out/r2e
O=gpurun_out/r2e
timeout 1200 python -m pytest tests -m gpu -q > $O/gputest.log 2>&1; echo "rc=$?" >> $O/gputest.log
for cfg in lm_ddp lm_fsdp; do
  timeout 300 python bench.py --config $cfg --steps 30 --warmup 8 > $O/${cfg}.json 2> $O/${cfg}.err
  HYPERION_CE_L2_ROWS=0 timeout 300 python bench.py --config $cfg --steps 30 --warmup 8 > $O/${cfg}_nol2.json 2> $O/${cfg}_nol2.err
  HYPERION_CE_L2_ROWS=1024 timeout 300 python bench.py --config $cfg --steps 30 --warmup 8 > $O/${cfg}_l2_1024.json 2> $O/${cfg}_l2_1024.err
  HYPERION_CE_L2_ROWS=256 timeout 300 python bench.py --config $cfg --steps 30 --warmup 8 > $O/${cfg}_l2_256.json 2> $O/${cfg}_l2_256.err
done
timeout 300 python bench.py --steps 10 --warmup 3 > $O/gpt2.json 2> $O/gpt2.err
HYPERION_CE_L2_ROWS=512 timeout 300 python bench.py --steps 10 --warmup 3 > $O/gpt2_l2_512.json 2> $O/gpt2_l2_512.err
timeout 300 python bench.py --config cifar_ddp --steps 30 --warmup 8 > $O/cifar.json 2> $O/cifar.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_lm.csv python bench.py --config lm_ddp --steps 2 --warmup 3 --no-graph --no-e2e > $O/ncu_lm.log 2>&1
python scripts/launch_summary.py $O/launches_lm.csv 30 adamw > $O/launches_lm_summary.txt 2>&1
tail -5 $O/gputest.log; for f in $O/*.json; do echo $f $(grep -o '"ms_per_step": [0-9.]*' $f | head -1); done; cat $O/launches_lm_summary.txt
