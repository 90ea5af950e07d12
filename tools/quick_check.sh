#!/bin/bash
# a subset of the GPU tests by -k expression + one short default bench line:  bash tools/quick_check.sh "<-k expr>" <tag>
o=gpurun_out; mkdir -p $o
python -m pytest tests -m gpu -x -q -k "$1" > $o/${2}_pytest.log 2>&1; echo "pytest rc $?"; tail -3 $o/${2}_pytest.log
python bench.py --steps 10 --warmup 3 --cpu-seconds 2 > $o/${2}_bench.json 2> $o/${2}_bench.err; echo "bench rc $?"
python - <<PY
import json
d = json.load(open("$o/${2}_bench.json"))
print("value %.1f (%.3f ms) e2e %.1f (%.3f ms) first %.3f nohint %.3f digest %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["first_frame_ms"], d["ms_per_step_no_hint"], d["parity"]["frame_digest"]))
PY
