#!/bin/bash
# usage: scripts/bench_matrix.sh N "cfg:impl[:flag...]" ...   -> gpurun_out/matrix/n<N>_<cfg>_<impl>[_flags].json
# One JSON line per run (bench.py contract); CUDA events, max over ranks; every arm carries a clocks record.
N=$1; shift
O=gpurun_out/matrix; mkdir -p $O
PORT=29540
for spec in "$@"; do
  IFS=: read -r cfg impl f1 f2 f3 <<< "$spec"
  flags=""; tag=""
  for f in $f1 $f2 $f3; do flags="$flags --$f"; tag="${tag}_$(echo $f | tr -d '= ' | tr '-' '_')"; done
  steps=20; warm=5
  case $cfg in llama_*) steps=8; warm=3;; lm_*|cifar_ddp) steps=40; warm=10;; esac
  out=$O/n${N}_${cfg}_${impl}${tag}
  PORT=$((PORT+1))
  if [ "$N" = "1" ]; then
    timeout 900 python bench.py --gpus 1 --config $cfg --impl $impl --steps $steps --warmup $warm $flags > $out.log 2> $out.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --config $cfg --impl $impl --steps $steps --warmup $warm $flags > $out.log 2> $out.err
  fi
  rc=$?
  grep -h '^{' $out.log > $out.json
  echo "$spec rc=$rc $(grep -o '"value": [0-9.]*' $out.json | head -1) $(grep -o '"ms_per_step": [0-9.]*' $out.json | head -1) $(grep -o '"unavailable": "[^"]\{0,80\}' $out.json)"
  [ $rc -ne 0 ] && tail -5 $out.err
done
