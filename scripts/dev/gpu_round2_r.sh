#!/bin/bash
mkdir -p gpurun_out/r2r
O=gpurun_out/r2r
timeout 600 python -m pytest tests -m gpu -q > $O/gputest.log 2>&1; echo "rc=$?" >> $O/gputest.log; tail -4 $O/gputest.log
for cfg in cifar_ddp resnet50_ddp lm_ddp; do
  timeout 300 python bench.py --config $cfg --steps 40 --warmup 10 > $O/$cfg.json 2> $O/$cfg.err; echo "$cfg $(grep -o '"ms_per_step": [0-9.]*' $O/$cfg.json | head -1) launches $(grep -o '"gpu_launches": [0-9]*' $O/$cfg.json)"
done
python -c "from __graft_entry__ import smoke; smoke()" 2>&1 | tail -1
