#!/bin/bash
# c5 (10M-triangle soup, 64-frame orbit) on 8 GPUs: gpurun --gpus 8 --timeout 900 -- 'bash tools/scale_c5.sh <tag>'
tag=${1:-c5}
o=gpurun_out; mkdir -p $o
timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29508 \
    bench.py --gpus 8 --workload c5 --steps 8 --warmup 3 > $o/${tag}_c5_n8.json 2> $o/${tag}_c5_n8.err; echo "rc $?"
python - <<PY
import json
d = json.load(open("$o/${tag}_c5_n8.json"))
print("value %.1f ms %.3f e2e %.1f (%.3f ms) first %.3f nohint %.3f digest %s identical %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["first_frame_ms"], d["ms_per_step_no_hint"], d["parity"]["frame_digest"], d["parity"]["pixels_identical"]))
PY
tail -2 $o/${tag}_c5_n8.err | cut -c1-200
