#!/bin/bash
# Round 2, GPU call 3: the restructured traversal (walk + two searches of the scene-wide hierarchy as functions of their own)
o=gpurun_out; t=r02c
python -m pytest tests -m gpu -x -q > $o/${t}_pytest.log 2>&1; echo "pytest rc $? : $(tail -1 $o/${t}_pytest.log)"
D=rust-ray-tracer_b200/csrc/librrt_b200.so
for m in 0 1 2; do echo "#### RRT_SEARCH_MODE=$m"; RRT_SEARCH_MODE=$m python tools/variant_sweep.py $D variants/mb6.so variants/mb5.so -- c2 c3a c3b c4; done > $o/${t}_modes.txt 2>&1
for m in 0 1; do echo "#### RRT_SEARCH_MODE=$m"; RRT_SEARCH_MODE=$m python tools/variant_sweep.py $D -- c5; done >> $o/${t}_modes.txt 2>&1
grep -E "^==|^####|ms" $o/${t}_modes.txt
RRT_LIB_PATH=$PWD/variants/probe.so python tools/probe.py c4 c2 > $o/${t}_probe.txt 2>&1; cat $o/${t}_probe.txt
python bench.py > $o/${t}_bench_c4.json 2> $o/${t}_bench_c4.err; echo "bench rc $?"; python -c "
import json; d=json.load(open('$o/${t}_bench_c4.json')); print(d['value'], d['ms_per_step'], d['e2e'])"
