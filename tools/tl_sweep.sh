#!/bin/bash
# development aid: drain profile for several RRT_ORDER_TAIL settings, then plain timings of the main library
for t in ${TAILS:-2 4 8}; do echo "== RRT_ORDER_TAIL=$t"; RRT_ORDER_TAIL=$t RRT_LIB_PATH=variants/tl/tl3.so python tools/drain_timeline.py c4 8 2>&1 | grep "hint 1"; done
for h in 0 1 0 1; do echo "== main library RRT_ORDER_HINT=$h"; RRT_ORDER_HINT=$h python tools/variant_sweep.py rust-ray-tracer_b200/csrc/librrt_b200.so -- ${WL:-c2 c4 c5}; done
