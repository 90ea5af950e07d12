#!/bin/bash
o=gpurun_out; mkdir -p $o
n=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
    tools/frame_parallel_vs_rows.py c4 16 > $o/fp_c4_n$n.json 2> $o/fp_c4_n$n.err; echo "rc $?"
cat $o/fp_c4_n$n.json; tail -5 $o/fp_c4_n$n.err | cut -c1-300
