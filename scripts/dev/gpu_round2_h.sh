#!/bin/bash
# final-tree check + ncu evidence (1 GPU; the only tool in this call is ncu)
mkdir -p gpurun_out/r2h
O=gpurun_out/r2h
timeout 900 python -m pytest tests -m gpu -q > $O/gputest.log 2>&1; echo "rc=$?" >> $O/gputest.log; tail -3 $O/gputest.log
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_gpt2.json 2> $O/bench_gpt2.err; echo "bench rc=$?"
timeout 300 python bench.py --config resnet50_ddp --steps 20 --warmup 5 > $O/bench_resnet50.json 2> $O/bench_resnet50.err
timeout 300 python bench.py --config lm_ddp --steps 20 --warmup 5 > $O/bench_lm_ddp.json 2> $O/bench_lm_ddp.err
python -c "from __graft_entry__ import smoke; smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"
timeout 300 python scripts/ncu_targets.py > $O/targets_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on --nvtx --nvtx-include "profile/" -f -o $O/prof python scripts/ncu_targets.py > $O/ncu.log 2>&1
echo "ncu rc=$?"; tail -3 $O/targets_plain.log; tail -3 $O/ncu.log; ls -la $O/*.ncu-rep
for f in $O/bench_*.json; do grep -o '"value": [0-9.]*\|"e2e": {"value": [0-9.]*\|"samples": [0-9]*\|"sm_mhz": [0-9.]*' $f | tr '\n' ' '; echo; done
