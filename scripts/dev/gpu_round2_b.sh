#!/bin/bash
# dropout kernels: numerics, then the whole GPU suite, then lm_ddp / lm_fsdp with native vs library dropout
mkdir -p gpurun_out/r2b
O=gpurun_out/r2b
timeout 600 python -m pytest tests/test_gpu_dropout.py -q -x > $O/dropout.log 2>&1; echo "rc=$?" >> $O/dropout.log
timeout 1200 python -m pytest tests -m gpu -q > $O/gputest.log 2>&1; echo "rc=$?" >> $O/gputest.log
for cfg in lm_ddp lm_fsdp; do
  timeout 300 python bench.py --config $cfg --steps 30 --warmup 8 > $O/${cfg}_native.json 2> $O/${cfg}_native.err
  HYPERION_DROPOUT=torch timeout 300 python bench.py --config $cfg --steps 30 --warmup 8 > $O/${cfg}_torchdrop.json 2> $O/${cfg}_torchdrop.err
done
timeout 300 python bench.py --steps 10 --warmup 3 > $O/gpt2.json 2> $O/gpt2.err
tail -15 $O/dropout.log; tail -5 $O/gputest.log; for f in $O/*.json; do echo $f; head -c 400 $f; echo; done
