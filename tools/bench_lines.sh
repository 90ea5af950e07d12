#!/bin/bash
# the three one-GPU bench lines of the default build (what profiles/r02_v11_bench_*_n1.json hold)
o=gpurun_out; mkdir -p $o
for w in c4 c2 c5; do
  python bench.py --workload $w > $o/line_$w.json 2> $o/line_$w.err; echo "bench $w rc $?"
done
