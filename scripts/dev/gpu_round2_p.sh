#!/bin/bash
mkdir -p gpurun_out/r2p
O=gpurun_out/r2p
timeout 600 python -m pytest tests -m gpu -q > $O/gputest.log 2>&1; echo "rc=$?" >> $O/gputest.log; tail -4 $O/gputest.log
timeout 300 python bench.py --steps 30 --warmup 8 > $O/gpt2.json 2> $O/gpt2.err; echo "gpt2 fused-colsum $(grep -o '"ms_per_step": [0-9.]*' $O/gpt2.json | head -1)"
HYPERION_GEMM_COLSUM=0 timeout 300 python bench.py --steps 30 --warmup 8 > $O/gpt2_nocs.json 2> $O/gpt2_nocs.err; echo "gpt2 separate colsum $(grep -o '"ms_per_step": [0-9.]*' $O/gpt2_nocs.json | head -1)"
timeout 300 python bench.py --config vit_ddp --steps 20 --warmup 5 > $O/vit.json 2> $O/vit.err; echo "vit $(grep -o '"ms_per_step": [0-9.]*' $O/vit.json | head -1)"
