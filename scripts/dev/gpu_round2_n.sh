#!/bin/bash
mkdir -p gpurun_out/r2n
O=gpurun_out/r2n
timeout 600 python -m pytest tests -m gpu -q > $O/gputest.log 2>&1; echo "rc=$?" >> $O/gputest.log; tail -3 $O/gputest.log
for v in "igemm igemm" "cudnn igemm" "igemm cudnn" "cudnn cudnn"; do set -- $v
  HYPERION_CONV_DGRAD=$1 HYPERION_CONV_WGRAD=$2 timeout 300 python bench.py --config resnet50_ddp --steps 30 --warmup 8 > $O/resnet50_d$1_w$2.json 2> $O/resnet50_d$1_w$2.err
  echo "resnet50 dgrad=$1 wgrad=$2 $(grep -o '"ms_per_step": [0-9.]*' $O/resnet50_d$1_w$2.json | head -1)"
  HYPERION_CONV_DGRAD=$1 HYPERION_CONV_WGRAD=$2 timeout 300 python bench.py --config cifar_ddp --steps 40 --warmup 10 > $O/cifar_d$1_w$2.json 2> $O/cifar_d$1_w$2.err
  echo "cifar dgrad=$1 wgrad=$2 $(grep -o '"ms_per_step": [0-9.]*' $O/cifar_d$1_w$2.json | head -1)"
done
timeout 200 python -m hyperion_b200 probe --out $O/results/hardware --sizes 1024 > $O/probe.log 2>&1; tail -8 $O/probe.log
