#!/bin/bash
# multi-GPU bench lines (run under gpurun --gpus 8, charged 8x):  bash tools/scale_run.sh <tag> [c5]
# Each line carries its own no-hint timing (ms_per_step_no_hint), first-frame timing and parity digest.
tag=${1:-scale}
o=gpurun_out
run() { # n workload timeout extra...
  n=$1; w=$2; to=$3; shift 3
  timeout $to python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --workload $w "$@" > $o/${tag}_${w}_n$n.json 2> $o/${tag}_${w}_n$n.err
  echo "$w n=$n rc $? $(python - <<PY
import json
try:
    d = json.load(open("$o/${tag}_${w}_n$n.json"))
    print("value %.1f ms %.3f e2e %.1f (%.3f ms) first %.3f nohint %.3f digest %s identical %s assembled==single %s" % (
        d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["first_frame_ms"], d["ms_per_step_no_hint"],
        d["parity"]["frame_digest"], d["parity"]["pixels_identical"], d["parity"]["assembled_equals_single_gpu_frame"]))
except Exception as e:
    print("no line:", e)
PY
)"
  tail -2 $o/${tag}_${w}_n$n.err | cut -c1-200
}
run 8 c4 240 --steps 20 --warmup 5
run 4 c4 240 --steps 20 --warmup 5
if [ "$2" = "c5" ]; then run 8 c5 700 --steps 8 --warmup 3; fi
