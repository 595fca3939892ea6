#!/bin/bash
# 2 GPUs: multi-GPU tests (incl. FULL_SHARD parity), then gpt2 / lm benches at N=2 in resident and reshard modes, reference arm at N=2
mkdir -p gpurun_out/r2c
O=gpurun_out/r2c
timeout 900 python -m pytest tests/test_gpu_multi.py -q -x > $O/multi.log 2>&1; echo "rc=$?" >> $O/multi.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 300 $TR bench.py --gpus 2 --steps 15 --warmup 4 > $O/gpt2_n2.json 2> $O/gpt2_n2.err
timeout 300 $TR bench.py --gpus 2 --steps 15 --warmup 4 --reshard > $O/gpt2_n2_reshard.json 2> $O/gpt2_n2_reshard.err
HYPERION_FSDP_AG=plain timeout 300 $TR bench.py --gpus 2 --steps 15 --warmup 4 --config lm_fsdp > $O/lm_fsdp_n2_plain.json 2> $O/lm_fsdp_n2_plain.err
for cfg in lm_ddp lm_fsdp cifar_ddp; do
  timeout 300 $TR bench.py --gpus 2 --steps 30 --warmup 8 --config $cfg > $O/${cfg}_n2.json 2> $O/${cfg}_n2.err
  timeout 300 $TR bench.py --gpus 2 --steps 30 --warmup 8 --config $cfg --impl reference > $O/${cfg}_n2_reference.json 2> $O/${cfg}_n2_reference.err
done
timeout 400 $TR bench.py --gpus 2 --steps 10 --warmup 3 --config llama_lora --size tiny > $O/llama_tiny_n2.json 2> $O/llama_tiny_n2.err
timeout 400 $TR bench.py --gpus 2 --steps 10 --warmup 3 --config llama_lora --size tiny --reshard > $O/llama_tiny_n2_reshard.json 2> $O/llama_tiny_n2_reshard.err
tail -12 $O/multi.log; for f in $O/*.json; do echo $f; grep '^{' $f | head -c 700; echo; done; tail -5 $O/*reshard.err
