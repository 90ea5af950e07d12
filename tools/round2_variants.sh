#!/bin/bash
# Builds the compile-time variants prepared at the end of round 1 (no GPU was left to time them) into variants/
# (git-ignored, travels with gpurun).  Every variant passes the host-emulation suite bit for bit:
#     RRT_EMU_FLAGS="<flags>" python -m pytest tests/test_host_emulation.py -q
# Build here, then on a GPU box:
#     gpurun --timeout 900 -- 'python tools/variant_sweep.py variants/*.so -- c2 c4 c5 > gpurun_out/variants.txt 2>&1'
# (variant_sweep.py checks that every variant renders the same frame and hit-id plane, sha1, and prints its time).
set -e
cd "$(dirname "$0")/.."
mkdir -p variants
build() {   # name flags...
  local name=$1; shift
  RRT_OUT=$PWD/variants/$name.so RRT_EXTRA_FLAGS="$*" python rust-ray-tracer_b200/build.py --force > /dev/null 2>&1
  printf "%-14s %-55s %s\n" "$name" "$*" "$(python tools/sass_count.py variants/$name.so --kernel render_kernelILb0ELb1 | grep -o '[0-9]* instructions')"
}
build a_default
build b_pred        -DRRT_SLAB_PRED=1
build c_l3p1        -DRRT_CULL_LOOP=3 -DRRT_PEND=1
build d_l3p2        -DRRT_CULL_LOOP=3 -DRRT_PEND=2
build e_l3p3        -DRRT_CULL_LOOP=3 -DRRT_PEND=3
build f_l3p4        -DRRT_CULL_LOOP=3 -DRRT_PEND=4
build g_pred_l3p2   -DRRT_SLAB_PRED=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2
build h_late        -DRRT_LATE_PRUNE=1
build i_pred_late   -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1
build j_all         -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2
build k_all_mb7     -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2 -DRRT_MIN_BLOCKS=7
build l_all_mb6     -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2 -DRRT_MIN_BLOCKS=6
build m_filter      -DRRT_SHADOW_FILTER=1
build n_filter_pred -DRRT_SHADOW_FILTER=1 -DRRT_SLAB_PRED=1
build o_filter_all  -DRRT_SHADOW_FILTER=1 -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2
build p_filter_prim -DRRT_SHADOW_FILTER=1 -DRRT_PRIMARY_FILTER=1
build q_filter_prim_all -DRRT_SHADOW_FILTER=1 -DRRT_PRIMARY_FILTER=1 -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2
# scene-wide culling hierarchy for the filter: run these with RRT_GLOBAL_CULL=1 (round2_first_call.sh does)
build r_gfilter     -DRRT_SHADOW_FILTER=2
build s_gfilter_prim -DRRT_SHADOW_FILTER=2 -DRRT_PRIMARY_FILTER=1
build t_gfilter_all -DRRT_SHADOW_FILTER=2 -DRRT_PRIMARY_FILTER=1 -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2
build u_sparse      -DRRT_SHADOW_FILTER=2 -DRRT_SPARSE_PRIMARY=1
build v_sparse_all  -DRRT_SHADOW_FILTER=2 -DRRT_SPARSE_PRIMARY=1 -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2
# register budget at today's code size (the 128 x 8 = 64-register choice dates from kernel v3)
build w_mb6         -DRRT_MIN_BLOCKS=6
build x_mb5         -DRRT_MIN_BLOCKS=5
build y_gfilter_mb6 -DRRT_SHADOW_FILTER=2 -DRRT_MIN_BLOCKS=6
build z_sparse_all_mb6 -DRRT_SHADOW_FILTER=2 -DRRT_SPARSE_PRIMARY=1 -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2 -DRRT_MIN_BLOCKS=6
build z_sparse_all_mb5 -DRRT_SHADOW_FILTER=2 -DRRT_SPARSE_PRIMARY=1 -DRRT_SLAB_PRED=1 -DRRT_LATE_PRUNE=1 -DRRT_CULL_LOOP=3 -DRRT_PEND=2 -DRRT_MIN_BLOCKS=5
