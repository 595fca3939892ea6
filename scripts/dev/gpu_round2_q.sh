#!/bin/bash
mkdir -p gpurun_out/r2q
O=gpurun_out/r2q
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29551 scripts/nvlink_bytes.py --config gpt2_fsdp > $O/nvlink_gpt2.log 2>&1
timeout 200 $TR --master-port 29552 scripts/nvlink_bytes.py --config gpt2_fsdp --reshard > $O/nvlink_gpt2_reshard.log 2>&1
timeout 200 $TR --master-port 29553 scripts/nvlink_bytes.py --config lm_ddp > $O/nvlink_lm_ddp.log 2>&1
grep -h NVLINK $O/*.log; tail -3 $O/nvlink_gpt2.log
