O=gpurun_out/r3i; mkdir -p $O
L=gaussian_splat_ipu_b200/libgsplat_b200.so
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > $O/tests.log; cat $O/tests.log
grep -q passed $O/tests.log || exit 1
bash profiles/ab_k5.sh $O/ab.txt sweep=$L three=$L:GSPLAT_SORT_SWEEP=0 sweep2=$L
