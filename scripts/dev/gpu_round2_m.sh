#!/bin/bash
# refresh of the committed result artefacts + A/B of the conv backward default + final-tree tests (1 GPU, no profiler)
mkdir -p gpurun_out/r2m
O=gpurun_out/r2m
timeout 600 python -m pytest tests -m gpu -q > $O/gputest.log 2>&1; echo "rc=$?" >> $O/gputest.log; tail -3 $O/gputest.log
for cfg in gpt2_fsdp lm_ddp lm_fsdp cifar_ddp resnet50_ddp; do
  timeout 300 python bench.py --config $cfg --steps 30 --warmup 8 > $O/bench_$cfg.json 2> $O/bench_$cfg.err; echo "$cfg rc=$? $(grep -o '"ms_per_step": [0-9.]*' $O/bench_$cfg.json | head -1)"
done
HYPERION_CONV_DGRAD=cudnn HYPERION_CONV_WGRAD=cudnn timeout 300 python bench.py --config resnet50_ddp --steps 30 --warmup 8 > $O/bench_resnet50_cudnnbwd.json 2> $O/bench_resnet50_cudnnbwd.err; echo "resnet50 cudnn-bwd $(grep -o '"ms_per_step": [0-9.]*' $O/bench_resnet50_cudnnbwd.json | head -1)"
timeout 300 python -m hyperion_b200 probe --out $O/results/hardware > $O/probe.log 2>&1; echo "probe rc=$?"
timeout 400 python -m hyperion_b200 baseline --out $O/results --precision fp32 --iters 20 --warmup 5 > $O/baseline_fp32.log 2>&1; echo "baseline fp32 rc=$?"; cp $O/results/baseline/model_benchmarks.csv $O/results/model_benchmarks_fp32.csv 2>/dev/null
timeout 400 python -m hyperion_b200 baseline --out $O/results --precision bf16 --iters 20 --warmup 5 > $O/baseline_bf16.log 2>&1; echo "baseline bf16 rc=$?"; cp $O/results/baseline/model_benchmarks.csv $O/results/model_benchmarks_bf16.csv 2>/dev/null
tail -20 $O/probe.log; cat $O/results/model_benchmarks_fp32.csv $O/results/model_benchmarks_bf16.csv
