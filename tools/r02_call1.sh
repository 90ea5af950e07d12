#!/bin/bash
# Round 2, GPU call 1: the GPU suite incl. the new full-size parity tests, the prepared kernel variants, the default
# bench line on this box, and the ncu capture of the kernel that SHIPS (render_kernel<false,true>, hint on).
#   gpurun --timeout 1500 -- 'bash tools/r02_call1.sh'
o=gpurun_out; t=r02a
python -m pytest tests -m gpu -x -q > $o/${t}_pytest.log 2>&1; echo "pytest rc $? : $(tail -1 $o/${t}_pytest.log)"
V=variants
python tools/variant_sweep.py $V/a_default.so $V/b_pred.so $V/d_l3p2.so $V/h_late.so $V/j_all.so $V/w_mb6.so $V/x_mb5.so -- c2 c3a c4 > $o/${t}_variants.txt 2>&1; echo "variants rc $?"
RRT_GLOBAL_CULL=1 python tools/variant_sweep.py $V/a_default.so $V/r_gfilter.so $V/y_gfilter_mb6.so $V/t_gfilter_all.so $V/u_sparse.so $V/v_sparse_all.so $V/z_sparse_all_mb6.so $V/z_sparse_all_mb5.so -- c2 c3a c4 > $o/${t}_variants_gcull.txt 2>&1; echo "gcull rc $?"
python tools/variant_sweep.py $V/a_default.so $V/j_all.so $V/w_mb6.so -- c5 > $o/${t}_variants_c5.txt 2>&1; echo "c5 rc $?"
grep -E "^==|ms" $o/${t}_variants.txt $o/${t}_variants_gcull.txt $o/${t}_variants_c5.txt
python bench.py > $o/${t}_bench_c4.json 2> $o/${t}_bench_c4.err; echo "bench rc $?"
python tools/profile_one.py c4 3 > $o/${t}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -f -o $o/${t}_c4 \
    python tools/profile_one.py c4 3 > $o/${t}_ncu_c4.log 2>&1; echo "ncu rc $?"
