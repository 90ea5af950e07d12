#!/bin/bash
# sweep (tools/sweep.sh) followed by one ncu --set full capture of the default library on c4
bash tools/sweep.sh "$1" "$2"
o=gpurun_out
ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 2 -c 1 -f -o $o/cur_c4 \
    python tools/profile_one.py c4 3 > $o/cur_ncu_c4.log 2>&1; echo "ncu c4 rc $?"
