#!/bin/bash
# evidence pass (1 GPU): layer GEMM bench vs cuBLAS, launch list of one GPT-2 step and one lm_ddp step (no graph), attention bench
mkdir -p gpurun_out/r2d
O=gpurun_out/r2d
timeout 300 python scripts/gemm_layer_bench.py > $O/gemm_layer.log 2>&1; cp gpurun_out/gemm_layer_bench.json $O/ 2>/dev/null
timeout 200 python scripts/attn_bench.py > $O/attn_bench.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_gpt2.csv python bench.py --steps 2 --warmup 3 --no-graph --no-e2e > $O/ncu_gpt2.log 2>&1
python scripts/launch_summary.py $O/launches_gpt2.csv 40 adamw > $O/launches_gpt2_summary.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_lm.csv python bench.py --config lm_ddp --steps 2 --warmup 3 --no-graph --no-e2e > $O/ncu_lm.log 2>&1
python scripts/launch_summary.py $O/launches_lm.csv 40 adamw > $O/launches_lm_summary.txt 2>&1
cat $O/gemm_layer.log; tail -20 $O/attn_bench.log; cat $O/launches_gpt2_summary.txt; cat $O/launches_lm_summary.txt
