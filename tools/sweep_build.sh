#!/bin/bash
# builds variants of the library with different launch bounds into gpurun_out-free temp names, then times them
# usage (on the GPU box): tools/sweep_build.sh "128:4 128:5 128:6 256:2 64:8"
set -e
cd "$(dirname "$0")/.."
for v in $1; do
  B=${v%%:*}; M=${v##*:}
  RRT_EXTRA_FLAGS="-DRRT_BLOCK=$B -DRRT_MIN_BLOCKS=$M" python rust-ray-tracer_b200/build.py --force > /dev/null 2>&1
  echo "== block $B minblocks $M: $(grep -A1 'render_kernelILb0' rust-ray-tracer_b200/csrc/build.log | grep -o 'Used [0-9]* registers' | head -1) $(grep -B1 'Used' rust-ray-tracer_b200/csrc/build.log | grep -m1 -o '[0-9]* bytes spill stores')"
  python tools/gpu_check.py model2 c4 2>&1 | grep -E "Mrays"
done
