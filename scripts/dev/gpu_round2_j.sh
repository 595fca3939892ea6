#!/bin/bash
# ncu evidence, part 2 (1 GPU): the backward kernels (launched from autograd's thread, outside the main thread's NVTX range)
mkdir -p gpurun_out/r2j
O=gpurun_out/r2j
timeout 300 python scripts/ncu_targets.py > $O/targets_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"flash_bwd|norm_bwd_kernel|conv_wgrad|bwd_dx_kernel|bwd_reduce_kernel|attn_delta|dq_convert" -c 30 -f -o /tmp/prof2 python scripts/ncu_targets.py > $O/ncu.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/prof2.ncu-rep --page raw --csv > $O/raw.csv 2> $O/raw.err
ncu -i /tmp/prof2.ncu-rep --page source --csv -k regex:flash_bwd64 -c 1 > $O/source_flash_bwd64_kernel.csv 2>> $O/raw.err
ls -la $O; tail -2 $O/ncu.log
