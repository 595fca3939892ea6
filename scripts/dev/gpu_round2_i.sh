#!/bin/bash
# ncu evidence (1 GPU): full-set profile of one launch per shipping kernel; pages exported to CSV on the box (the .ncu-rep itself
# is larger than what a call can bring back)
mkdir -p gpurun_out/r2i
O=gpurun_out/r2i
timeout 300 python scripts/ncu_targets.py > $O/targets_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on --nvtx --nvtx-include "profile/" -f -o /tmp/prof python scripts/ncu_targets.py > $O/ncu.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/prof.ncu-rep --page raw --csv > $O/raw.csv 2> $O/raw.err
ncu -i /tmp/prof.ncu-rep --page details --csv > $O/details.csv 2>> $O/raw.err
for k in gemm2_kernel flash_bwd64_kernel flash_fwd_kernel norm_bwd_kernel ce_fwd_bwd_reg_kernel conv_fwd_kernel; do
  ncu -i /tmp/prof.ncu-rep --page source --csv -k regex:$k > $O/source_$k.csv 2>> $O/raw.err
done
ls -la $O; du -sh $O
