#!/bin/bash
# under gpurun --gpus 2: the tests that need two GPUs (skipped on a one-GPU box) and a 2-GPU bench line
#   gpurun --gpus 2 --timeout 900 -- 'bash tools/two_gpu_check.sh r02t'
tag=${1:-two}
o=gpurun_out; mkdir -p $o
python -m pytest tests -m gpu -x -q -k "two_gpus or native_driver or sharded or peer" > $o/${tag}_pytest2.log 2>&1; echo "pytest rc $?"; tail -3 $o/${tag}_pytest2.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 \
    bench.py --gpus 2 --steps 10 --warmup 3 > $o/${tag}_c4_n2.json 2> $o/${tag}_c4_n2.err; echo "bench n2 rc $?"
python - <<PY
import json
d = json.load(open("$o/${tag}_c4_n2.json"))
print("value %.1f ms %.3f e2e %.1f first %.3f nohint %.3f digest %s assembled==single %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["first_frame_ms"], d["ms_per_step_no_hint"], d["parity"]["frame_digest"], d["parity"]["assembled_equals_single_gpu_frame"]))
PY
