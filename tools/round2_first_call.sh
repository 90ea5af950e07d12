#!/bin/bash
# First GPU call of round 2 (everything written after round 1's GPU budget ended, in one go):
#   here  :  bash tools/round2_variants.sh            (builds variants/*.so, ~6 min of CPU)
#   then  :  gpurun --timeout 1500 -- 'bash tools/round2_first_call.sh'
# 1. the GPU test suite incl. the two tests never run on a GPU (rrt_render_frames, frame-parallel world 1) and smoke()
# 2. every prepared kernel variant on c2 / c3a / c4 / c5: time + sha1 of frame and hit-id plane (must equal a_default's)
# 3. the default bench line (c4), so that the variant times have this box's baseline next to them
o=gpurun_out
python -m pytest tests -m gpu -x -q > $o/r02_pytest.log 2>&1; echo "pytest rc $? : $(tail -1 $o/r02_pytest.log)"
python __graft_entry__.py smoke > $o/r02_smoke.log 2>&1; echo "smoke rc $? : $(tail -1 $o/r02_smoke.log)"
python tools/variant_sweep.py variants/*.so -- c2 c3a c4 c5 > $o/r02_variants.txt 2>&1; echo "variants rc $?"
# the filter builds twice more (same libraries, same frames): filter always on (RRT_SHADOW_FILTER_RT=2: its pure effect;
# the sweep above ran the frame-to-frame switch, which turns it off on c5) and never on (=0: what the second loop costs by being there)
for m in 2 0; do
  RRT_SHADOW_FILTER_RT=$m python tools/variant_sweep.py variants/m_filter.so variants/o_filter_all.so -- c2 c4 c5 > $o/r02_variants_filter_rt$m.txt 2>&1; echo "filter rt=$m rc $?"
done
# the filter over ONE scene-wide culling hierarchy (flatten builds it when RRT_GLOBAL_CULL=1): expected to be the fastest form
RRT_GLOBAL_CULL=1 python tools/variant_sweep.py variants/a_default.so variants/r_gfilter.so variants/s_gfilter_prim.so variants/t_gfilter_all.so variants/u_sparse.so variants/v_sparse_all.so -- c2 c3a c4 > $o/r02_variants_gfilter.txt 2>&1; echo "gfilter rc $?"
grep -E "^==|ms" $o/r02_variants.txt $o/r02_variants_gfilter.txt | tail -120
# the whole GPU parity suite once more THROUGH the build with every switch on (rrt_trace_rays / rrt_ray_colours / frames
# against the oracle: the GPU-side proof of the switches, beyond the sha1 comparison of the sweep)
RRT_GLOBAL_CULL=1 RRT_LIB_PATH=$PWD/variants/v_sparse_all.so python -m pytest tests -m gpu -x -q > $o/r02_pytest_sparse_all.log 2>&1; echo "pytest (all switches) rc $? : $(tail -1 $o/r02_pytest_sparse_all.log)"
python bench.py > $o/r02_bench_c4_default.json 2> $o/r02_bench_c4_default.err; echo "bench rc $?"
