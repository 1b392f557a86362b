O=gpurun_out/r3a; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/tests.log; cat $O/tests.log
L=gaussian_splat_ipu_b200
bash profiles/ab_k5.sh $O/ab.txt old=$L/libgsplat_b200_old.so new=$L/libgsplat_b200.so w100=$L/libgsplat_b200_w100.so w400=$L/libgsplat_b200_w400.so h2k=$L/libgsplat_b200_h2k.so nopred=$L/libgsplat_b200_nopred.so nohl=$L/libgsplat_b200_nohl.so nohoist=$L/libgsplat_b200_nohoist.so old2=$L/libgsplat_b200_old.so
