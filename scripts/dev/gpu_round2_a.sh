#!/bin/bash
# First GPU pass of round 2 (1 GPU): tests, the headline bench, the reference-arm configs on all three arms.
mkdir -p gpurun_out/r2a
O=gpurun_out/r2a
timeout 1200 python -m pytest tests -m gpu -x -q > $O/gputest.log 2>&1; echo "pytest rc=$?" >> $O/gputest.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_gpt2.json 2> $O/bench_gpt2.err; echo "rc=$?" >> $O/bench_gpt2.err
for cfg in lm_ddp lm_fsdp cifar_ddp; do
  for impl in reference torch hyperion; do
    timeout 600 python bench.py --impl $impl --config $cfg --steps 30 --warmup 8 > $O/bench_${cfg}_${impl}.json 2> $O/bench_${cfg}_${impl}.err; echo "rc=$?" >> $O/bench_${cfg}_${impl}.err
  done
done
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > $O/bench_gpt2_reference.json 2>&1
timeout 900 python bench.py --impl torch --steps 10 --warmup 3 > $O/bench_gpt2_torch.json 2> $O/bench_gpt2_torch.err
timeout 300 torchrun --standalone --local-addr 127.0.0.1 --nproc_per_node=1 run_distributed.py --model gpt2 --steps 20 --epochs 2 --base_dir /tmp/rd > $O/cli_gpt2.log 2>&1
tail -3 $O/gputest.log; cat $O/bench_gpt2.json; for f in $O/bench_*_*.json; do echo $f; head -c 600 $f; echo; done; tail -3 $O/cli_gpt2.log
