#!/bin/bash
mkdir -p gpurun_out/r2o
O=gpurun_out/r2o
timeout 600 python -m pytest tests -m gpu -q > $O/gputest.log 2>&1; echo "rc=$?" >> $O/gputest.log; tail -3 $O/gputest.log
for cfg in lm_ddp gpt2_fsdp; do
  HYPERION_CE_BLOCK=512 timeout 300 python bench.py --config $cfg --steps 30 --warmup 8 > $O/${cfg}_ce512.json 2> $O/${cfg}_ce512.err
  echo "$cfg ce512 $(grep -o '"ms_per_step": [0-9.]*' $O/${cfg}_ce512.json | head -1)"
done
scripts/bench_matrix.sh 1 gpt2_fsdp:hyperion lm_ddp:hyperion lm_fsdp:hyperion cifar_ddp:hyperion resnet50_ddp:hyperion vit_ddp:hyperion llama_lora:hyperion llama_fsdp:hyperion
