#!/bin/bash
# direct host stores (RRT_HOST_STORES=1, default) against render + copy (=0): tests that read frames through rrt_render,
# then the e2e figure of the bench line both ways
o=gpurun_out; mkdir -p $o
python -m pytest tests -m gpu -x -q -k "parity or properties or no_alloc or native" > $o/hs_pytest.log 2>&1; echo "pytest rc $?"; tail -2 $o/hs_pytest.log
for hs in 1 0 1 0; do
  for w in c4 c2; do
    RRT_HOST_STORES=$hs python bench.py --workload $w --steps 10 --warmup 3 --cpu-seconds 2 > $o/hs_${w}_$hs.json 2> $o/hs_${w}_$hs.err
    python - <<PY
import json
d = json.load(open("$o/hs_${w}_$hs.json"))
print("host_stores=$hs $w value %.1f (%.3f ms) e2e %.1f (%.3f ms) digest %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["parity"]["frame_digest"]))
PY
  done
done
