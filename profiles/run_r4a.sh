O=gpurun_out/r4a; mkdir -p $O
L=gaussian_splat_ipu_b200/libgsplat_b200.so
bash profiles/ab_k5.sh $O/ab.txt default=$L split4=$L:GSPLAT_K5_SPLIT=4 later4=$L:GSPLAT_K5_SPLIT_LATER=4 later1=$L:GSPLAT_K5_SPLIT_LATER=1 s4l4=$L:"GSPLAT_K5_SPLIT=4 GSPLAT_K5_SPLIT_LATER=4"
