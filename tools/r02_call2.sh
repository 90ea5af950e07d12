#!/bin/bash
# Round 2, GPU call 2: standalone cost of each traversal flavour (tools/probe.py)
o=gpurun_out; t=r02b
for v in probe_default probe_default_mb8; do echo "== $v"; RRT_LIB_PATH=$PWD/variants/$v.so python tools/probe.py c4 c2 c3a; done > $o/${t}_probe.txt 2>&1
for v in probe_sparse probe_sparse_mb6 probe_sparse_mb8; do echo "== $v (RRT_GLOBAL_CULL=1)"; RRT_GLOBAL_CULL=1 RRT_LIB_PATH=$PWD/variants/$v.so python tools/probe.py c4 c2 c3a; done >> $o/${t}_probe.txt 2>&1
cat $o/${t}_probe.txt
