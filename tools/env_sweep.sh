#!/bin/bash
# usage (GPU box): tools/env_sweep.sh "RRT_CULL_THRESHOLD=8 RRT_CULL_LEAF=4" "RRT_CULL_THRESHOLD=2 RRT_CULL_LEAF=2" ...
cd "$(dirname "$0")/.."
for e in "$@"; do
  echo "== $e"
  env $e python tools/variant_sweep.py --one c2 c4
done
