O=gpurun_out/r4b; mkdir -p $O
L=gaussian_splat_ipu_b200/libgsplat_b200.so
bash profiles/ab_k5.sh $O/ab.txt three=$L sweep444=$L:GSPLAT_SORT_SWEEP=1 sweep296=$L:"GSPLAT_SORT_SWEEP=1 GSPLAT_SORT_SWEEP_GRID=296" sweep148=$L:"GSPLAT_SORT_SWEEP=1 GSPLAT_SORT_SWEEP_GRID=148" sweep74=$L:"GSPLAT_SORT_SWEEP=1 GSPLAT_SORT_SWEEP_GRID=74"
