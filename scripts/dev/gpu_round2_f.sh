#!/bin/bash
# 2 GPUs: block-unit all-gather GEMM: parity tests + GPT-2 A/B (fused into c_attn GEMM vs stand-alone gathers), lm configs
mkdir -p gpurun_out/r2f
O=gpurun_out/r2f
timeout 900 python -m pytest tests/test_gpu_multi.py -q -x > $O/multi.log 2>&1; echo "rc=$?" >> $O/multi.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 300 $TR bench.py --gpus 2 --steps 15 --warmup 4 > $O/gpt2_n2_agfused.json 2> $O/gpt2_n2_agfused.err
HYPERION_FSDP_AG=plain timeout 300 $TR bench.py --gpus 2 --steps 15 --warmup 4 > $O/gpt2_n2_agplain.json 2> $O/gpt2_n2_agplain.err
timeout 300 $TR bench.py --gpus 2 --steps 15 --warmup 4 --reshard > $O/gpt2_n2_reshard.json 2> $O/gpt2_n2_reshard.err
for cfg in lm_ddp lm_fsdp; do
  timeout 300 $TR bench.py --gpus 2 --steps 30 --warmup 8 --config $cfg > $O/${cfg}_n2.json 2> $O/${cfg}_n2.err
done
tail -12 $O/multi.log; for f in $O/*.json; do echo $f $(grep -o '"ms_per_step": [0-9.]*' $f | head -1) $(grep -o '"comm": "[^"]*"' $f); done; tail -3 $O/gpt2_n2_agfused.err
