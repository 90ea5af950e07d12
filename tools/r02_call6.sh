#!/bin/bash
# Round 2, GPU call 6: GPU suite on the restructured traversal, the adaptive switch, the new bench line, ncu of what ships
o=gpurun_out; t=r02f
python -m pytest tests -m gpu -x -q > $o/${t}_pytest.log 2>&1; echo "pytest rc $? : $(tail -1 $o/${t}_pytest.log)"
RRT_MODES="both,adaptive" python tools/search_modes.py c2 c3a c3b c4 c5 > $o/${t}_modes.txt 2>&1; cat $o/${t}_modes.txt
python bench.py > $o/${t}_bench_c4.json 2> $o/${t}_bench_c4.err; echo "bench rc $?"; tail -3 $o/${t}_bench_c4.err
python bench.py --workload c2 --cpu-seconds 8 > $o/${t}_bench_c2.json 2> $o/${t}_bench_c2.err; echo "bench c2 rc $?"
python tools/profile_one.py c4 3 > $o/${t}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 2 -c 1 -f -o $o/${t}_c4 \
    python tools/profile_one.py c4 3 > $o/${t}_ncu_c4.log 2>&1; echo "ncu rc $?"
