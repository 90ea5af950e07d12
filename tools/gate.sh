#!/bin/bash
# Round gate on the GPU box: tests, smoke, bench lines (both arms), ncu launch list and one --set full capture per workload.
#   gpurun --timeout 1500 -- 'bash tools/gate.sh r13'
tag=${1:-gate}
o=gpurun_out
python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest rc $?"
python __graft_entry__.py smoke > $o/${tag}_smoke.log 2>&1; echo "smoke rc $?"
for w in c4 c2 c5; do
  python bench.py --workload $w > $o/${tag}_bench_$w.json 2> $o/${tag}_bench_$w.err; echo "bench $w rc $?"
done
python bench.py --impl reference > $o/${tag}_bench_ref.json 2> $o/${tag}_bench_ref.err; echo "ref rc $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 > $o/${tag}_launches.log 2>&1; echo "launch list rc $?"
for w in c4 c2 c5; do
  ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -f -o $o/${tag}_$w \
      python tools/profile_one.py $w 3 > $o/${tag}_ncu_$w.log 2>&1; echo "ncu $w rc $?"
done
