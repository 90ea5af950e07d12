#!/bin/bash
# multi-GPU bench lines (run under gpurun --gpus 8):  bash tools/scale_run.sh <tag>
tag=${1:-scale}
o=gpurun_out
run() { # n workload extra...
  n=$1; w=$2; shift 2
  timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --workload $w "$@" > $o/${tag}_${w}_n$n$SUF.json 2> $o/${tag}_${w}_n$n$SUF.err
  echo "$w n=$n rc $? $(cut -c1-160 $o/${tag}_${w}_n$n$SUF.json)"
}
run 8 c4
run 8 c5 --steps 16 --warmup 3
run 4 c4
run 2 c4
SUF=_nohint RRT_ORDER_HINT=0 run 8 c4
