#!/bin/bash
o=gpurun_out; t=r02h
python -m pytest tests -m gpu -x -q > $o/${t}_pytest.log 2>&1; echo "pytest rc $? : $(tail -1 $o/${t}_pytest.log)"; grep -E "FAILED|Error|error" $o/${t}_pytest.log | head
for r in wave mega; do echo "== RRT_RENDERER=$r"; RRT_RENDERER=$r RRT_MODES="walk only,both,adaptive" python tools/search_modes.py c2 c3a c3b c4 c5; done > $o/${t}_modes.txt 2>&1; cat $o/${t}_modes.txt
