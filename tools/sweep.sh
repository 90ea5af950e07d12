#!/bin/bash
# Times every variants/*.so (tools/variant_sweep.py) on the workloads given ("c2 c4" by default); $2: libraries that
# also run c5.   gpurun --timeout 1200 -- 'bash tools/sweep.sh "c2 c3a c3b c4" "variants/a.so variants/b.so"'
o=gpurun_out
mkdir -p $o
python tools/variant_sweep.py variants/*.so -- ${1:-c2 c4} > $o/sweep.txt 2>&1
[ -n "$2" ] && python tools/variant_sweep.py $2 -- c5 > $o/sweep_c5.txt 2>&1
cat $o/sweep.txt; [ -n "$2" ] && cat $o/sweep_c5.txt
