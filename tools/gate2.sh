#!/bin/bash
# Round-2 gate on one GPU: tests, smoke, bench lines (c4 default, c2, c5), the reference arm, the ncu launch list of the
# bench command and one ncu --set full capture per workload (3rd launch), all into gpurun_out/<tag>_*.
#   gpurun --timeout 2400 -- 'bash tools/gate2.sh r02h'
tag=${1:-gate}
o=gpurun_out
mkdir -p $o
python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest rc $?"; tail -3 $o/${tag}_pytest.log
python __graft_entry__.py smoke > $o/${tag}_smoke.log 2>&1; echo "smoke rc $?"
for w in c4 c2 c5; do
  python bench.py --workload $w > $o/${tag}_bench_$w.json 2> $o/${tag}_bench_$w.err; echo "bench $w rc $?"
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 > $o/${tag}_launches.log 2>&1; echo "launch list rc $?"
for w in c4 c2 c5; do
  ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 2 -c 1 -f -o $o/${tag}_$w \
      python tools/profile_one.py $w 3 > $o/${tag}_ncu_$w.log 2>&1; echo "ncu $w rc $?"
done
python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_ref.json 2> $o/${tag}_bench_ref.err; echo "ref rc $?"
